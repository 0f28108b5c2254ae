"""Async SGD / FTRL throughput on CTR-shaped data (BASELINE configs 3 and 4).

    python benchmarks/async_sgd_bench.py --rows 400000 --minibatch 20000 --tau 4 [--quant]
    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/async_sgd_bench.py ...

One process per GPU (worker + server shard).  The first data pass localizes and caches every
minibatch (key dictionary, remap, table slots); the timed region is a later pass over the cached
minibatches: one fused worker kernel (pull -> SpMV/loss/grad -> push over NVLink) and the server
apply per step, CUDA events, max over ranks.  Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=400000, help="examples per GPU")
    ap.add_argument("--groups", type=int, default=8)
    ap.add_argument("--keys-per-group", type=int, default=1 << 20)
    ap.add_argument("--minibatch", type=int, default=20000)
    ap.add_argument("--tau", type=int, default=4, help="max_delay")
    ap.add_argument("--quant", action="store_true", help="int8 block-scaled push (fixing_float)")
    ap.add_argument("--algo", default="FTRL")
    ap.add_argument("--passes", type=int, default=3, help="timed passes after the caching pass")
    ap.add_argument("--debug", action="store_true", help="print every step, dump stacks after 25 s")
    args = ap.parse_args()
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
    from parameter_server_b200.parallel.comm import SingleComm, TorchComm
    from parameter_server_b200.utils.synthetic import SparseCSR, make_ctr_like

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        comm = TorchComm()
    else:
        comm = SingleComm()

    groups = make_ctr_like(args.rows, args.groups, args.keys_per_group, seed=11, shard=rank, device=dev)
    # one CSR over all groups (keys already carry the group id in their top bits)
    n = args.rows
    rows = torch.cat([torch.repeat_interleave(torch.arange(n, device=dev), g.row_ptr[1:] - g.row_ptr[:-1])
                      for g in groups.values()])
    keys = torch.cat([g.keys for g in groups.values()])
    # spread the hashed ids over the whole 64-bit key space so that the even split of Range::all()
    # (postmaster.cc:17-31) gives every server shard the same share (odd multiplier: a bijection)
    keys = keys * (0x9E3779B97F4A7C15 - (1 << 64))
    order = torch.argsort(rows, stable=True)
    rp = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    rp[1:] = torch.cumsum(torch.bincount(rows, minlength=n), 0)
    data = SparseCSR(rp, keys[order].contiguous(), None, next(iter(groups.values())).y, 1 << 63)
    nnz = int(keys.numel())
    del groups, rows, keys, order

    # open addressing with linear probing: keep the load factor of a shard's table below 1/2
    cap = 1
    while cap < 2 * args.groups * args.keys_per_group // max(world, 1):
        cap <<= 1
    cfg = AsyncSGDConfig(algo=args.algo, minibatch=args.minibatch, max_delay=args.tau, alpha=0.1, beta=1.0,
                         lambda_l1=1.0, lambda_l2=0.0, fixing_float_by_nbytes=1 if args.quant else 0,
                         table_capacity=max(cap, 1 << 20), report_every=1 << 30)
    eng = AsyncSGDEngine(cfg, comm, dev, out=open(os.devnull, "w"))
    eng.load(data)

    if args.debug:
        import faulthandler
        faulthandler.dump_traceback_later(25, exit=True)

    def one_pass():
        seen = 0
        for i in range(eng.steps_per_pass):
            m = eng.step(i if i < eng.n_minibatches else eng.n_minibatches)
            seen += m.n_rows
            if args.debug and i % 5 == 4:
                print(f"rank {rank} enqueued step {i} seq {eng.seq}", file=sys.stderr, flush=True)
            if args.debug and os.environ.get("PSB_DEBUG_SYNC"):
                torch.cuda.synchronize()
                print(f"rank {rank} step {i} seq {eng.seq} flags {eng.b_lflag.tensor.tolist()} "
                      f"acks {eng.b_ack.tensor.tolist()}", file=sys.stderr, flush=True)
        return seen

    one_pass()                      # localize + cache every minibatch, fill the tables
    one_pass()                      # warm-up on the cached path
    torch.cuda.synchronize()
    comm.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = eng.launches
    e0.record()
    seen = 0
    for _ in range(args.passes):
        seen += one_pass()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(seen)], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    st = eng.wstats.cpu().tolist()
    if rank == 0:
        steps = args.passes * eng.steps_per_pass
        print(json.dumps({
            "metric": "examples/sec, async SGD (%s) on CTR-shaped hashed binary features" % args.algo,
            "value": tot.item() / (ms.item() / 1e3), "unit": "examples/s", "n_gpus": world,
            "ms_per_minibatch_step": ms.item() / steps, "steps": steps,
            "config": {"rows_per_gpu": args.rows, "nnz_per_gpu": nnz, "minibatch": args.minibatch,
                       "max_delay": args.tau, "push": "int8 block-scaled" if args.quant else "fp32",
                       "key_space": args.groups * args.keys_per_group, "table_capacity_per_shard": cfg.table_capacity,
                       "timing": "CUDA events over cached-minibatch passes, max over ranks",
                       "mean_logloss_last_window": st[0] / max(st[2], 1.0)},
            "gpu_launches": eng.launches - l0}))
    eng.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
