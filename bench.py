#!/usr/bin/env python
"""Headline benchmark: Darlin (L1 logistic regression, block coordinate descent) on
rcv1-shaped synthetic data - examples/sec over whole passes, one rank per GPU.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun)
    python bench.py --impl reference ...                     (reference arm)
    python bench.py --impl nccl ...                          (our engine with NCCL collectives
                                                              instead of the peer-memory kernels)

A "step" is one full pass of Darlin over the training matrix (all feature blocks:
gradient+push, server prox update, pull+dual update, plus the progress evaluation), i.e.
what the reference's scheduler does per iteration (darlin.h:58-103).  Scaling is weak: every
rank holds one rcv1-sized shard (677 399 x 47 236), N=1 is exactly BASELINE.json's config.

Phases of a run (everything after the first is outside the timed region unless stated):
  1. load + preprocess (load_preprocess_s), pass graph captured
  2. W warm-up passes + K TIMED passes from the zero model          -> value, e2e
  3. invariant check with plain fp64 torch: margins == y * X w from the gathered shards,
     objective recomputed == objective reported                      -> verified
  4. time to objective: back to the zero model, the reference's rule (epsilon 1e-4 twice in a
     row with a filter reset, at most max_pass_of_data = 20 passes)   -> config.time_to_objective
  5. N > 1: strong scaling - the ONE 677 399-row problem split over the N ranks, K timed
     passes + time to objective                                      -> config.strong
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def reference_arm(args) -> None:
    """The unmodified reference, if it could be installed under baseline/_ref (DESIGN.md section 5
    records why it cannot on this image): nothing of this repository runs on that path."""
    exe = os.path.join(ROOT, "baseline", "_ref", "build", "ps")
    if not os.path.exists(exe):
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is a make-built C++ tree (no setup.py/pyproject) whose generated "
                          "*.pb.cc are not in the tree and need protoc 2.5 + libprotobuf, zeromq, gflags, "
                          "glog, snappy headers; none on the image, no network (DESIGN.md section 5)"}))
        return
    print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref/build/ps present but its "
                      "third-party runtime libraries are not loadable on this image"}))


class Runner:
    """The reference's scheduler loop (darlin.h:52-120) around DarlinEngine.run_pass."""

    def __init__(self, eng, epsilon: float):
        self.eng, self.eps = eng, epsilon
        self.restart()

    def restart(self):
        self.kkt, self.reset, self.objs, self.wall, self.prog = 1e20, False, [], [], []
        self.converged_at = None

    def one_pass(self, it: int):
        eng = self.eng
        t = time.perf_counter()
        order, n_prior = eng.next_order(it)
        p = eng.run_pass(order, self.kkt, self.reset, n_prior)  # public API: H2D + pass + D2H (host sync)
        self.wall.append(time.perf_counter() - t)
        self.kkt = p.violation / max(eng.num_ex_global, 1) * eng.cfg.kkt_filter_threshold_ratio
        rel = self.objs[-1] / p.objective - 1 if self.objs else 1.0
        self.objs.append(p.objective)
        self.prog.append(p)
        if 0 < rel <= self.eps:
            if self.reset and self.converged_at is None:
                self.converged_at = it
            self.reset = True
        else:
            self.reset = False
        return p


def time_to_objective(eng, args, dist, dev, world) -> dict:
    """Phase 4: the reference's stopping rule from the zero model."""
    import torch
    eng.reset_model()
    run = Runner(eng, args.epsilon)
    it = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    while True:
        go = run.converged_at is None and it < args.max_pass
        if world > 1:  # every rank sees the same objective; a collective decision cannot diverge
            flag = torch.tensor([int(go)], dtype=torch.int64, device=dev)
            dist.broadcast(flag, 0)
            go = bool(flag.item())
        if not go:
            break
        run.one_pass(it)
        it += 1
    ev1.record()
    torch.cuda.synchronize()
    t = torch.tensor([sum(run.wall), ev0.elapsed_time(ev1) / 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    last = run.prog[-1]
    return {"seconds": t[0].item(), "device_span_seconds": t[1].item(), "passes": it,
            "objective": run.objs[-1], "nnz_w": last.nnz_w, "nnz_active_set": last.nnz_active_set,
            "converged": run.converged_at is not None,
            "hit_pass_cap": run.converged_at is None and it >= args.max_pass,
            "rule": "relative objective change <= %g on two consecutive passes, the second after a KKT "
                    "filter reset (darlin.h:103-118), at most %d passes (rcv1/batch_l1lr.conf); wall "
                    "clock of every pass from the zero model incl. H2D, D2H, host sync, max over ranks"
                    % (args.epsilon, args.max_pass)}


def timed_passes(eng, args, comm, dist, dev, world, rank, local, sample_clocks: bool) -> dict:
    """Phase 2: W warm-up + K timed passes from the zero model."""
    import torch
    run = Runner(eng, 0.0)

    def barrier():
        comm.barrier()
        torch.cuda.synchronize()

    for it in range(args.warmup):
        run.one_pass(it)
    barrier()
    from parameter_server_b200.utils.clocks import ClockSampler
    sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
    if sampler:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev_ms = ker_ms = 0.0
    t_wall0 = time.perf_counter()
    ev0.record()
    for it in range(args.warmup, args.warmup + args.steps):
        p = run.one_pass(it)
        dev_ms += p.device_ms
        ker_ms += p.kernel_ms
    ev1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    span_ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if sampler else None
    times = torch.tensor([dev_ms, span_ms, t_wall * 1e3, ker_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms, span_ms, wall_ms, ker_ms = times.tolist()
    first, last = run.prog[args.warmup], run.prog[-1]
    return {"dev_ms": dev_ms, "span_ms": span_ms, "wall_ms": wall_ms, "ker_ms": ker_ms, "clocks": clocks,
            "objs": run.objs, "last_objective": run.objs[-1],
            "active_first_last": [first.nnz_active_set, last.nnz_active_set],
            "nnz_w_first_last": [first.nnz_w, last.nnz_w]}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl"])
    ap.add_argument("--config", default="darlin", choices=["darlin", "ctr_sgd"],
                    help="darlin: the headline (BASELINE configs 1 / 2); ctr_sgd: BASELINE configs 3 / 4, streaming "
                         "async SGD at the named CTR size - delegates to benchmarks/ctr_sgd_stream.py, whose own "
                         "options (--examples-per-gpu, --passes, --push ...) may follow; it is timed from the first "
                         "minibatch of the first pass, so --steps / --warmup do not apply")
    ap.add_argument("--rows", type=int, default=677399, help="examples per GPU")
    ap.add_argument("--features", type=int, default=47236)
    ap.add_argument("--nnz-per-row", type=float, default=73.2)
    ap.add_argument("--tau", type=int, default=0, help="max_block_delay")
    ap.add_argument("--epsilon", type=float, default=1e-4, help="stopping rule (rcv1/batch_l1lr.conf)")
    ap.add_argument("--max-pass", type=int, default=20, help="max_pass_of_data (rcv1/batch_l1lr.conf)")
    ap.add_argument("--block-ratio", type=float, default=4.0)
    ap.add_argument("--seg-len", type=int, default=128)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--profile-phases", action="store_true")
    ap.add_argument("--fused-grid", type=int, default=0)
    ap.add_argument("--split", action="store_true", help="3 kernels per step instead of the persistent pass kernel")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-benchmark (N > 1)")
    ap.add_argument("--no-tto", action="store_true", help="skip the time-to-objective phase")
    args, rest = ap.parse_known_args()
    if args.config == "ctr_sgd" and args.impl == "ours":
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "benchmarks"))
        import ctr_sgd_stream
        ctr_sgd_stream.main(rest)
        return
    if rest:
        ap.error("unrecognized arguments: " + " ".join(rest))
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:
            reference_arm(args)
        return

    import torch
    from parameter_server_b200.models.darlin import DarlinConfig, DarlinEngine
    from parameter_server_b200.parallel.comm import SingleComm, TorchComm
    from parameter_server_b200.utils.synthetic import make_rcv1_like

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or (world == 1 and args.gpus == 1), \
        f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist, t_nccl = None, 0.0
    if world > 1:
        import torch.distributed as dist
        t0 = time.time()
        dist.init_process_group("nccl", device_id=dev)
        # communicator bootstrap (channels of every collective used during set-up) is not preprocessing
        dist.all_reduce(torch.zeros(1, device=dev))
        dist.all_to_all_single(torch.zeros(world, device=dev), torch.zeros(world, device=dev))
        dist.all_gather([torch.zeros(1, device=dev) for _ in range(world)], torch.zeros(1, device=dev))
        dist.all_gather_object([None] * world, rank)
        dist.barrier()
        torch.cuda.synchronize()
        t_nccl = time.time() - t0
        comm = TorchComm()
    else:
        comm = SingleComm()

    nccl_arm = args.impl == "nccl"
    cfg = DarlinConfig(lambda_l1=1.0, eta=1.0, max_pass_of_data=args.max_pass,
                       epsilon=args.epsilon, feature_block_ratio=args.block_ratio,
                       max_block_delay=args.tau, seg_len=args.seg_len, use_graph=not args.no_graph,
                       fused=not (args.split or nccl_arm), collectives="nccl" if nccl_arm else "peer",
                       profile_phases=args.profile_phases, fused_grid=args.fused_grid)

    def build(data):
        t0 = time.time()
        eng = DarlinEngine(cfg, comm, dev, out=sys.stderr)
        eng.load(data)
        eng.capture()
        torch.cuda.synchronize()
        return eng, time.time() - t0

    t0 = time.time()
    data = make_rcv1_like(args.rows, args.features, args.nnz_per_row, seed=0, shard=rank, device=dev)
    t_gen = time.time() - t0
    eng, t_load = build(data)
    nnz = eng.nnz
    del data

    res = timed_passes(eng, args, comm, dist, dev, world, rank, local, True)
    check = eng.verify(res["last_objective"])
    if args.profile_phases and rank == 0:
        pr = eng.prof.cpu().tolist()
        passes = args.steps + args.warmup
        nst = max(len(eng.blocks) // eng.bps, 1)
        names = ['phase1', 'plan+prefetch', 'barrierA', 'phase3', 'barrierC']
        print('per-step us (mean over CTAs):', {n: round(v / eng.grid_fused / passes / nst / 1e3, 2)
                                                for n, v in zip(names, pr)}, file=sys.stderr)
    tto = None if args.no_tto else time_to_objective(eng, args, dist, dev, world)
    n_ex = eng.num_ex_global
    n_steps_pass = len(eng.plan_steps(list(eng.blk_order)))
    launches, h2d, d2h = eng.launches_per_pass, eng.h2d_bytes_per_pass, eng.d2h_bytes_per_pass
    n_blocks, grid = len(eng.blocks), eng.grid_fused
    eng.close()
    del eng
    torch.cuda.empty_cache()

    strong = None
    if world > 1 and not args.no_strong:
        # the ONE rcv1-sized problem of N=1 (shard 0), rows dealt out in contiguous slices
        full = make_rcv1_like(args.rows, args.features, args.nnz_per_row, seed=0, shard=0, device=dev)
        lo, hi = rank * args.rows // world, (rank + 1) * args.rows // world
        a, b = int(full.row_ptr[lo]), int(full.row_ptr[hi])
        from parameter_server_b200.utils.synthetic import SparseCSR
        part = SparseCSR((full.row_ptr[lo:hi + 1] - a).contiguous(), full.keys[a:b].contiguous(),
                         None if full.vals is None else full.vals[a:b].contiguous(),
                         full.y[lo:hi].contiguous(), full.n_features)
        del full
        eng2, t_load2 = build(part)
        del part
        r2 = timed_passes(eng2, args, comm, dist, dev, world, rank, local, False)
        chk2 = eng2.verify(r2["last_objective"])
        tto2 = None if args.no_tto else time_to_objective(eng2, args, dist, dev, world)
        strong = {"global_examples": eng2.num_ex_global, "ms_per_step": r2["ker_ms"] / args.steps,
                  "examples_per_s": eng2.num_ex_global * args.steps / (r2["ker_ms"] / 1e3),
                  "us_per_block_step": r2["ker_ms"] / args.steps * 1e3 / max(len(eng2.plan_steps(list(eng2.blk_order))), 1),
                  "verified": chk2["verified"], "load_preprocess_s": t_load2, "time_to_objective": tto2}
        eng2.close()

    if rank == 0:
        ker_ms, wall_ms = res["ker_ms"], res["wall_ms"]
        value = n_ex * args.steps / (ker_ms / 1e3)
        e2e = n_ex * args.steps / (wall_ms / 1e3)
        kernel = ("NCCL collectives (all_to_all + fixed-order reduce, all_gather) between split kernels" if nccl_arm
                  else "split (3 launches/step)" if args.split else "persistent cooperative pass kernel")
        out = {
            "metric": "examples/sec, L1-LR Darlin BCD (one step = one full pass over the data)",
            "value": value, "unit": "examples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ker_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp64 state / fp32 matrix values",
            "data": "synthetic rcv1-shaped (seeded), zero-init weights",
            "impl": args.impl, "verified": check["verified"],
            "config": {"model": "L1 logistic regression, Darlin block coordinate descent",
                       "examples_per_gpu": args.rows, "global_examples": n_ex,
                       "features": args.features, "nnz_per_gpu": nnz,
                       "feature_blocks": n_blocks, "max_block_delay": args.tau,
                       "block_steps_per_pass": n_steps_pass,
                       "us_per_block_step": ker_ms / args.steps * 1e3 / max(n_steps_pass, 1),
                       "kernel": kernel, "grid": grid,
                       "lambda": 1.0, "parallelism": f"dp{world} workers x {world} key-range server shards",
                       "l2": "inputs larger than L2 (CSC index+value arrays %.0f MB per GPU)" % (nnz * 8 / 1e6),
                       "timing": "value: CUDA events around the kernels of each pass (graph replay), summed, max over ranks; e2e: wall clock incl. host scheduler, pinned H2D, D2H, sync",
                       "timed_region": "passes %d..%d from the zero model; KKT filter running as in the reference" % (args.warmup, args.warmup + args.steps - 1),
                       "nnz_active_set_first_last_timed": res["active_first_last"],
                       "nnz_w_first_last_timed": res["nnz_w_first_last"],
                       "ms_per_step_with_copies": res["dev_ms"] / args.steps,
                       "objective_first_last": [res["objs"][0], res["objs"][-1]],
                       "verification": check,
                       "load_preprocess_s": t_load, "generate_s": t_gen, "nccl_bootstrap_s": t_nccl,
                       "time_to_objective": tto, "strong": strong},
            "clocks": res["clocks"],
            "e2e": {"value": e2e, "unit": "examples/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "wall_ms_per_step": wall_ms / args.steps,
                    "api": "DarlinEngine.run_pass (pinned order+ctrl H2D, graph replay, progress D2H, host sync) per step"},
            "gpu_launches": launches * args.steps,
            "event_span_ms": res["span_ms"],
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
