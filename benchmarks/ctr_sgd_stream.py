"""BASELINE configs 3 and 4: CTR-shaped L1-LR by asynchronous SGD (FTRL) with bounded delay,
hyper-parameters of example/linear/ctr/online_l1lr.conf (minibatch 10 000, 10 data passes,
max_delay 4, tail_feature_freq 1, lambda (10, 1), DECAY alpha .01 beta 10), at the named size:
100 M examples over 1 B hashed feature ids on 8 GPUs (12.5 M examples per GPU).

    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/ctr_sgd_stream.py
    python benchmarks/ctr_sgd_stream.py --examples-per-gpu 2000000 --passes 3       (one GPU, smaller)

The examples live in PINNED HOST memory (what the reference's StreamReader + prefetch queue hold,
src/learner/sgd.h:104-157); every pass streams them again: one async H2D copy per array and
minibatch into a ring of device buffers, localisation (sort / unique / CountMin tail filter / remap
/ table slots; one enqueue, no host sync) on the copy stream, all issued by the native runner
(csrc/cuda/sgd_stream.cu), then per step the fused worker kernel (pull over NVLink -> margins,
loss, gradient -> push, fp32 / int8 / fp8) and the owner's apply.  Nothing is cached: the first
pass and the later ones do the same work, except that later passes find their keys in the tables.
Timed from the first minibatch of the first pass with CUDA events + wall clock, max over ranks.
Prints one JSON line (bench.py schema) on rank 0.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def make_chunk(n_rows, n_groups, ids_per_group, nnz_per_group, seed, dev):
    """CTR-shaped binary examples generated on the device: per feature group a few ids per row,
    popularity ~ rank^(-3/4), ids hashed into 54 bits with the group in the top 10 (TERAFEA layout,
    src/data/text_parser.cc:129-181), then spread over the whole 64-bit key space.  The label comes
    from a planted sparse model whose weight is a hash of the key (no table of 1 B weights)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    M54 = (1 << 54) - 1
    rows, keys = [], []
    margin = torch.zeros(n_rows, dtype=torch.float32, device=dev)
    for grp in range(1, n_groups + 1):
        ln = torch.poisson(torch.full((n_rows,), float(nnz_per_group), device=dev), generator=g).clamp_(min=1).long()
        row = torch.repeat_interleave(torch.arange(n_rows, device=dev), ln)
        u = torch.rand(row.numel(), generator=g, device=dev, dtype=torch.float64)
        rank = (u ** 4 * ids_per_group).to(torch.int64).clamp_(max=ids_per_group - 1)
        h = (rank * (0x9E3779B97F4A7C15 - (1 << 64)) + ((grp * 0x632BE59BD9B4E019) & M54)) & M54
        key = h | (grp << 54)
        # planted weight: 2 % of the ids matter, value in [-2, 2) from other bits of the hash
        hw = (key * (0xD6E8FEB86659FD93 - (1 << 64)))
        active = ((hw >> 20) & 63) == 0
        wv = (((hw >> 32) & 0xFFFF).float() / 16384.0 - 2.0) * active
        margin.index_add_(0, row, wv)
        rows.append(row)
        keys.append(key * (0x9E3779B97F4A7C15 - (1 << 64)))   # odd multiplier: a bijection of the key space
    row = torch.cat(rows)
    key = torch.cat(keys)
    order = torch.argsort(row, stable=True)
    counts = torch.bincount(row, minlength=n_rows)
    noise = torch.randn(n_rows, generator=g, device=dev)
    y = torch.where(margin - 1.0 + 0.7 * noise > 0, 1.0, -1.0)   # CTR-like: positives are the minority
    return counts, key[order], y


class HostData:
    """One worker's examples in pinned host memory, cut into minibatches ahead of time (each with
    its own row offsets) so that feeding a minibatch is pure pointer arithmetic."""

    def __init__(self, n_rows, minibatch, n_groups, ids_per_group, nnz_per_group, seed, dev):
        from parameter_server_b200.utils.synthetic import SparseCSR
        self.SparseCSR = SparseCSR
        self.mb = minibatch
        self.n_rows = n_rows
        est = int(n_rows * n_groups * (nnz_per_group + 0.2) * 1.05) + 1024
        self.keys = torch.empty(est, dtype=torch.int64).pin_memory()
        self.y = torch.empty(n_rows, dtype=torch.float32).pin_memory()
        self.n_mb = (n_rows + minibatch - 1) // minibatch
        self.row_ptr = torch.zeros(self.n_mb, minibatch + 1, dtype=torch.int64).pin_memory()
        self.mb_lo = [0] * (self.n_mb + 1)
        chunk = minibatch * max(1, 500_000 // minibatch)
        at_row = at_key = 0
        while at_row < n_rows:
            n = min(chunk, n_rows - at_row)
            counts, keys, y = make_chunk(n, n_groups, ids_per_group, nnz_per_group, seed * 100003 + at_row, dev)
            assert at_key + keys.numel() <= est
            self.keys[at_key:at_key + keys.numel()].copy_(keys)
            self.y[at_row:at_row + n].copy_(y)
            counts = counts.cpu()
            for b in range(0, n, minibatch):
                i = (at_row + b) // minibatch
                c = counts[b:b + minibatch]
                self.row_ptr[i, 1:c.numel() + 1] = torch.cumsum(c, 0)
                self.row_ptr[i, c.numel() + 1:] = self.row_ptr[i, c.numel()]
                self.mb_lo[i + 1] = self.mb_lo[i] + int(c.sum())
            at_row += n
            at_key += keys.numel()
        self.nnz = at_key
        self.max_nnz_mb = max(self.mb_lo[i + 1] - self.mb_lo[i] for i in range(self.n_mb))

    def batches(self):
        for i in range(self.n_mb):
            a = i * self.mb
            rows = min(self.mb, self.n_rows - a)
            yield self.SparseCSR(self.row_ptr[i, :rows + 1], self.keys[self.mb_lo[i]:self.mb_lo[i + 1]], None,
                                 self.y[a:a + rows], 1 << 63)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--examples-per-gpu", type=int, default=12_500_000)
    ap.add_argument("--passes", type=int, default=10)
    ap.add_argument("--minibatch", type=int, default=10_000)
    ap.add_argument("--groups", type=int, default=8)
    ap.add_argument("--ids-per-group", type=int, default=1 << 27, help="8 x 2^27 = 1.07 B hashed ids")
    ap.add_argument("--nnz-per-group", type=float, default=3.0)
    ap.add_argument("--tau", type=int, default=4)
    ap.add_argument("--tail-freq", type=int, default=1)
    ap.add_argument("--push", default="fp32", choices=["fp32", "int8", "fp8"])
    ap.add_argument("--table-capacity", type=int, default=0, help="slots per shard (0: from the data size)")
    args = ap.parse_args(argv)
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
    from parameter_server_b200.ops.evaluation import auc as dev_auc
    from parameter_server_b200.parallel.comm import SingleComm, TorchComm

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        comm = TorchComm()
    else:
        comm = SingleComm()

    t0 = time.time()
    data = HostData(args.examples_per_gpu, args.minibatch, args.groups, args.ids_per_group, args.nnz_per_group,
                    seed=11 + rank, dev=dev)
    t_gen = time.time() - t0
    # distinct keys a shard can see: bounded by the id space and by what the data can touch; open
    # addressing wants the load below 1/2
    cap = args.table_capacity
    if cap <= 0:
        touch = min(args.groups * args.ids_per_group, int(0.35 * data.nnz * world)) // world
        cap = 1 << 20
        while cap < 2 * touch:
            cap <<= 1
    cfg = AsyncSGDConfig(algo="FTRL", loss="LOGIT", minibatch=args.minibatch, num_data_pass=args.passes,
                         max_delay=args.tau, alpha=0.01, beta=10.0, lambda_l1=10.0, lambda_l2=1.0,
                         lr_type="DECAY", tail_feature_freq=args.tail_freq, countmin_n=100_000_000, countmin_k=2,
                         fixing_float_by_nbytes=0 if args.push == "fp32" else 1,
                         push_format="fp8" if args.push == "fp8" else "int8", table_capacity=cap,
                         report_every=1 << 30, prefetch=4)
    eng = AsyncSGDEngine(cfg, comm, dev, out=open(os.devnull, "w"))
    eng.open(data.max_nnz_mb + 128)
    steps = int(max(comm.gather_obj(data.n_mb)))
    torch.cuda.synchronize()
    comm.barrier()

    per_pass = []
    last = {}

    def keep_last(i, m):
        if i == steps - 1 and m.n_rows:
            last["y"], last["xw"] = m.y.clone(), m.xw[:m.n_rows].clone()

    from parameter_server_b200.utils.clocks import ClockSampler
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.passes + 1)]
    l0 = eng.launches
    wall0 = time.perf_counter()
    ev[0].record()
    total = 0
    for p in range(args.passes):
        seen = eng.train_stream(data.batches(), steps, on_step=keep_last)
        ev[p + 1].record()
        ev[p + 1].synchronize()                      # D2H of the pass statistics (the step's result)
        st = eng.wstats.cpu().tolist()
        eng.wstats.zero_()
        eng.check_fault()
        total += seen
        per_pass.append({"pass": p, "ms": ev[p].elapsed_time(ev[p + 1]), "logloss": st[0] / max(st[2], 1.0),
                         "accuracy": st[1] / max(st[2], 1.0),
                         "auc_last_minibatch": dev_auc(last["y"], last["xw"]) if last else float("nan")})
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ev[0].elapsed_time(ev[-1]), wall * 1e3, per_pass[0]["ms"],
                      sum(pp["ms"] for pp in per_pass[1:]) / max(len(per_pass) - 1, 1)], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(total), float(data.nnz)], dtype=torch.float64, device=dev)
    load = torch.tensor([eng.check_table_load()], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
        dist.all_reduce(load, op=dist.ReduceOp.MAX)
    if os.environ.get("PSB_SGD_TIMING") and eng._runner is not None:
        print(f"rank {rank}: stage ms/step {eng._runner.timing()}", file=sys.stderr, flush=True)
    if os.environ.get("PSB_LOC_TIMING"):
        import ctypes
        out7 = (ctypes.c_double * 7)()
        eng.lib.psb_sgd_localize_timing.restype = ctypes.c_int64
        calls = eng.lib.psb_sgd_localize_timing(out7)
        names = ("sort", "rle", "sketch", "select", "columns", "owner_begin", "slots")
        print(f"rank {rank}: localise ms/call over {calls} calls "
              + str({k: round(v, 4) for k, v in zip(names, out7)}), file=sys.stderr, flush=True)
    if rank == 0:
        ms, wall_ms, first_ms, steady_ms = t.tolist()
        ex_pass = tot[0].item() / args.passes
        print(json.dumps({
            "metric": "examples/sec, async SGD (FTRL, bounded delay) on CTR-shaped hashed binary features, streamed from pinned host memory",
            "value": tot[0].item() / (ms / 1e3), "unit": "examples/s", "n_gpus": world, "steps": steps * args.passes,
            "warmup": 0, "ms_per_step": ms / (steps * args.passes), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32 weights / optimizer state, %s push" % args.push,
            "data": "synthetic CTR-shaped (seeded), zero-init model",
            "config": {"model": "L1+L2 logistic regression, async SGD FTRL (example/linear/ctr/online_l1lr.conf)",
                       "global_examples": int(ex_pass), "examples_per_gpu": args.examples_per_gpu,
                       "hashed_id_space": args.groups * args.ids_per_group, "nnz_global": int(tot[1].item()),
                       "minibatch": args.minibatch, "passes": args.passes, "max_delay": args.tau,
                       "tail_feature_freq": args.tail_freq, "table_capacity_per_shard": cap,
                       "table_load_max": load.item(), "push": args.push,
                       "first_pass_examples_per_s": ex_pass / (first_ms / 1e3),
                       "steady_examples_per_s": ex_pass / (steady_ms / 1e3) if args.passes > 1 else None,
                       "per_pass": per_pass, "generate_s": t_gen,
                       "rank0_host_enqueue_s": eng.submit_s, "rank0_host_source_s": eng.source_s,
                       "host_cpus": os.cpu_count(),
                       "timing": "CUDA events from the first minibatch of pass 0 (nothing cached or warmed), max over ranks"},
            "e2e": {"value": tot[0].item() / (wall_ms / 1e3), "unit": "examples/s",
                    "h2d_bytes_per_step": eng.h2d_bytes / max(steps * args.passes, 1), "d2h_bytes_per_step": 32.0 / steps,
                    "api": "AsyncSGDEngine.train_stream"},
            "gpu_launches": eng.launches - l0, "clocks": clocks}))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
