"""What resolving keys in a REMOTE open-addressing table costs (psb_kv_lookup_insert: the probing
code the async-SGD localisation uses), separated by case: keys that are new (a system-scope CAS
over NVLink per key) against keys that are present (reads only), one rank probing against all ranks
probing one another at once, and the same against the local shard.

    torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/peer_table_probe.py
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from parameter_server_b200._lib import check, current_stream, ptr
from parameter_server_b200.models.async_sgd import _lib
from parameter_server_b200.parallel.comm import TorchComm


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", torch.cuda.current_device())
    backend = os.environ.get("PROBE_BACKEND", "nccl")   # gloo: no NCCL communicator in the process at all
    if backend == "nccl":
        dist.init_process_group("nccl", device_id=dev)
    else:
        dist.init_process_group("gloo")
    comm, lib = TorchComm(), _lib()
    cap = 1 << 28
    arena = comm.make_arena(cap * 8 + 4096, dev)
    b_keys = arena.alloc("tab_keys", cap, torch.int64)
    check(lib.psb_kv_table_clear(b_keys.ptr, cap, current_stream()))
    torch.cuda.synchronize()
    comm.barrier()
    P = C.c_void_p
    tabs = (P * world)(*[arena.peer_ptr(s, b_keys) for s in range(world)])
    caps = (C.c_int64 * world)(*[cap] * world)
    overflow = torch.zeros(1, dtype=torch.int32, device=dev)
    n = 200_000
    g = torch.Generator(device=dev).manual_seed(17 + rank)
    results = {}

    def run(name, keys, owner, active):
        """time one lookup_insert call of `n` keys on the ranks in `active` (the others idle)"""
        slots = torch.empty(n, dtype=torch.int64, device=dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        comm.barrier()
        if rank in active:
            a.record()
            check(lib.psb_kv_lookup_insert(ptr(keys), n, ptr(owner), tabs, caps, world, ptr(slots), ptr(overflow),
                                           current_stream()))
            b.record()
        torch.cuda.synchronize()
        comm.barrier()
        t = torch.tensor([a.elapsed_time(b) * 1e3 if rank in active else 0.0], dtype=torch.float64,
                         device=dev if backend == "nccl" else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        results[name] = round(t.item(), 1)

    # warm-up: load the kernel (lazy module loading) outside the timed calls
    one = torch.zeros(1, dtype=torch.int64, device=dev)
    check(lib.psb_kv_lookup_insert(ptr(one + 5 + rank), 1, ptr(torch.full((1,), rank, dtype=torch.int32, device=dev)),
                                   tabs, caps, world, ptr(one.clone()), ptr(overflow), current_stream()))
    everyone = set(range(world))
    for where, own in (("local", rank), ("peer", (rank + 1) % world)):
        owner = torch.full((n,), own, dtype=torch.int32, device=dev)
        for who, active in (("one rank", {0}), ("all ranks", everyone)):
            keys = torch.randint(0, 1 << 62, (n,), generator=g, device=dev, dtype=torch.int64)
            run(f"{where}, {who}, new keys (CAS)", keys, owner, active)
            run(f"{where}, {who}, present keys (reads)", keys, owner, active)
    if rank == 0:
        print(json.dumps({"what": "microseconds per psb_kv_lookup_insert call", "keys_per_call": n, "n_gpus": world, "backend": backend,
                          "env": {k: v for k, v in os.environ.items() if k.startswith("NCCL_")},
                          "table_slots": cap, "us": results}, indent=1))
    arena.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
