#!/usr/bin/env python
"""Push / pull bandwidth sweep (BASELINE.json config 5): n sparse keys per rank against a
key-range-sharded device KV vector, peer-memory kernels vs an NCCL all-to-all variant.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        benchmarks/push_pull_sweep.py [--sizes 1e3,1e4,...] [--density 0.5] [--nccl]

Every rank pushes / pulls the values of `n` sorted keys drawn from a dictionary of n/density
keys.  Bytes counted per rank and direction: n * 4 (fp32 payload; keys never travel after
register()).  Time: CUDA events, flushing nothing (payloads above ~32 MB exceed L2 anyway), max
over ranks, median of `--iters` runs after 3 warm-ups.  Roofline: 770 GB/s measured peer copy per
direction (B200_PROFILING.md); traffic that stays on the local GPU (1/N of it) is counted too.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from parameter_server_b200.parallel.comm import SingleComm, TorchComm
from parameter_server_b200.parallel.kv import DeviceKVVector


def timed(fn, iters, sync):
    ts = []
    for i in range(iters + 3):
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1e3,1e4,1e5,1e6,1e7,1e8")
    ap.add_argument("--density", type=float, default=0.5)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--nccl", action="store_true", help="also time the NCCL all-to-all variant")
    ap.add_argument("--tcp", action="store_true", help="also time the host TCP van on the same payload (rank 0)")
    ap.add_argument("--check-up-to", type=float, default=2e7, help="verify the final state for sizes up to this")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        comm = TorchComm()
    else:
        comm = SingleComm()
    rows = []
    for n in [int(float(x)) for x in a.sizes.split(",")]:
        n_dict = int(n / a.density)
        g = torch.Generator(device=dev).manual_seed(7)
        dictionary = torch.arange(n_dict, device=dev, dtype=torch.int64) * 3 + 11  # sorted, sparse key space
        kv = DeviceKVVector(dictionary, comm, dev, bulk_capacity=n)
        gsel = torch.Generator(device=dev).manual_seed(100 + rank)
        sel = torch.randperm(n_dict, generator=gsel, device=dev)[:n].sort().values
        h = kv.register(dictionary[sel], bulk=True)
        vals = torch.randn(max(n, 4), generator=g, device=dev)[:n].contiguous()
        out = torch.empty(max(n, 4), dtype=torch.float32, device=dev)

        def sync():
            kv.fence()
            torch.cuda.synchronize()

        t_push = timed(lambda: kv.push(h, vals), a.iters, sync)
        t_pull = timed(lambda: kv.pull(h, out), a.iters, sync)
        # bulk push, zero copy: the application writes its values into the registered staging view
        # (as one does with a registered NCCL buffer); bulk_push_copy_ms includes a D2D copy into it
        stage = kv.push_buffer(h)
        stage.copy_(vals)
        t_bpush = timed(lambda: kv.push_bulk(h, stage), a.iters, sync)
        t_bpush_copy = timed(lambda: kv.push_bulk(h, vals), a.iters, sync)
        t_bpull = timed(lambda: kv.pull_bulk(h), a.iters, sync)
        rec = {"n_keys": n, "push_ms": t_push, "pull_ms": t_pull, "bulk_push_ms": t_bpush,
               "bulk_push_copy_ms": t_bpush_copy, "bulk_pull_ms": t_bpull}
        # correctness of the final state: every rank pushed `vals` at its own key selection
        # 3 * (iters + 3) times (direct, bulk, bulk with copy); a bulk pull must return exactly that sum
        if n <= a.check_up_to:
            sync()
            n_push = 3 * (a.iters + 3)
            want = torch.zeros(n, dtype=torch.float64, device=dev)
            for r in range(world):
                if r == rank:
                    want += vals.double()
                    continue
                gr = torch.Generator(device=dev).manual_seed(100 + r)
                other = torch.randperm(n_dict, generator=gr, device=dev)[:n].sort().values
                pos = torch.searchsorted(other, sel).clamp_(max=n - 1)
                hit = other[pos] == sel
                want += torch.where(hit, vals[pos].double(), torch.zeros((), dtype=torch.float64, device=dev))
                del other, pos, hit
            got = kv.pull_bulk(h).double()
            err = ((got - n_push * want).abs() / (1 + (n_push * want).abs())).max().item()
            direct = ((kv.pull(h, out).double() - n_push * want).abs() / (1 + (n_push * want).abs())).max().item()
            rec["max_rel_error"] = max(err, direct)
            rec["correct"] = bool(max(err, direct) < 1e-4)
            assert rec["correct"], rec
        if a.tcp and rank == 0 and n <= 10_000_000:
            # the CPU plumbing path: the same payload (keys + values) through the framed-TCP van
            # between two nodes on this host, 4 hops (ref van_test.cc ring)
            import time
            from parameter_server_b200.runtime import host
            t0 = time.perf_counter()
            host.van_ring(2, True, 21000 + (os.getpid() + len(rows) * 7) % 9000, n)
            rec["tcp_van_ms_per_hop"] = (time.perf_counter() - t0) * 1e3 / 4
            rec["tcp_van_gbs"] = n * 12 / (rec["tcp_van_ms_per_hop"] * 1e-3) / 1e9
        if a.nccl and world > 1:
            # NCCL variant: values sorted by owner -> all_to_all_single -> local index_add / gather back
            counts = torch.bincount(h.owner[:n].long(), minlength=world)
            in_split = counts.tolist()
            cnt_out = torch.empty_like(counts)
            dist.all_to_all_single(cnt_out, counts)
            out_split = cnt_out.tolist()
            slots_recv = torch.empty(sum(out_split), dtype=torch.int64, device=dev)
            dist.all_to_all_single(slots_recv, h.slot[:n].contiguous(), out_split, in_split)
            recv = torch.empty(sum(out_split), dtype=torch.float32, device=dev)
            shard = kv.value

            def nccl_push():
                dist.all_to_all_single(recv, vals, out_split, in_split)
                shard.index_add_(0, slots_recv, recv)

            back = torch.empty(n, dtype=torch.float32, device=dev)

            def nccl_pull():
                dist.all_to_all_single(back, shard[slots_recv], in_split, out_split)

            rec["nccl_push_ms"] = timed(nccl_push, a.iters, lambda: (dist.barrier(), torch.cuda.synchronize()))
            rec["nccl_pull_ms"] = timed(nccl_pull, a.iters, lambda: (dist.barrier(), torch.cuda.synchronize()))
        t = torch.tensor([rec[k] for k in sorted(rec) if k.endswith("_ms")], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        for k, v in zip([k for k in sorted(rec) if k.endswith("_ms")], t.tolist()):
            rec[k] = v
        gbs = lambda ms: n * 4 / (ms * 1e-3) / 1e9  # noqa: E731
        rec.update(bulk_push_gbs=gbs(rec["bulk_push_ms"]), bulk_pull_gbs=gbs(rec["bulk_pull_ms"]),
                   bulk_push_copy_gbs=gbs(rec["bulk_push_copy_ms"]),
                   bulk_push_frac_of_770=gbs(rec["bulk_push_ms"]) / 770,
                   bulk_pull_frac_of_770=gbs(rec["bulk_pull_ms"]) / 770)
        rec.update(push_gbs=gbs(rec["push_ms"]), pull_gbs=gbs(rec["pull_ms"]),
                   push_frac_of_770=gbs(rec["push_ms"]) / 770, pull_frac_of_770=gbs(rec["pull_ms"]) / 770)
        if "nccl_push_ms" in rec:
            rec.update(nccl_push_gbs=gbs(rec["nccl_push_ms"]), nccl_pull_gbs=gbs(rec["nccl_pull_ms"]))
        rows.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        kv.close()
    if rank == 0:
        print(json.dumps({"n_gpus": world, "density": a.density, "sweep": rows}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
