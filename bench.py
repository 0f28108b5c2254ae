#!/usr/bin/env python
"""Headline benchmark: Darlin (L1 logistic regression, block coordinate descent) on
rcv1-shaped synthetic data — examples/sec over whole passes, one rank per GPU.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun)
    python bench.py --impl reference ...                     (reference arm)

A "step" is one full pass of Darlin over the training matrix (all feature blocks:
gradient+push, server prox update, pull+dual update, plus the progress evaluation), i.e.
what the reference's scheduler does per iteration (darlin.h:58-103).  Scaling is weak: every
rank holds one rcv1-sized shard (677 399 x 47 236), N=1 is exactly BASELINE.json's config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def reference_arm(args) -> None:
    # The reference is a make-built C++ tree (no setup.py / pyproject), and its third-party
    # stack (protobuf 2.5 + protoc, ZeroMQ, gflags, glog, snappy, Eigen) is not on the image.
    ok = os.path.isdir(os.path.join(ROOT, "baseline", "_ref")) and \
        os.path.exists(os.path.join(ROOT, "baseline", "_ref", "build", "ps"))
    if not ok:
        print(json.dumps({"impl": "reference", "unavailable":
                          "pip install of /root/reference fails (no setup.py/pyproject; C++ make "
                          "tree needing protobuf-2.5/zeromq/gflags/glog/snappy, none on the image, no network)"}))
        return
    print(json.dumps({"impl": "reference", "unavailable": "reference binary present but no runner"}))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--rows", type=int, default=677399, help="examples per GPU")
    ap.add_argument("--features", type=int, default=47236)
    ap.add_argument("--nnz-per-row", type=float, default=73.2)
    ap.add_argument("--tau", type=int, default=0, help="max_block_delay")
    ap.add_argument("--epsilon", type=float, default=1e-4, help="stopping rule (rcv1/batch_l1lr.conf)")
    ap.add_argument("--max-pass", type=int, default=20, help="max_pass_of_data (rcv1/batch_l1lr.conf)")
    ap.add_argument("--block-ratio", type=float, default=4.0)
    ap.add_argument("--seg-len", type=int, default=256)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--profile-phases", action="store_true")
    ap.add_argument("--kernel-flags", type=int, default=0)
    ap.add_argument("--fused-grid", type=int, default=0)
    ap.add_argument("--split", action="store_true", help="3 kernels per step instead of the persistent pass kernel")
    args = ap.parse_args()
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
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        comm = TorchComm()
    else:
        comm = SingleComm()

    cfg = DarlinConfig(lambda_l1=1.0, eta=1.0, max_pass_of_data=args.steps + args.warmup,
                       epsilon=0.0, feature_block_ratio=args.block_ratio,
                       max_block_delay=args.tau, seg_len=args.seg_len, use_graph=not args.no_graph, fused=not args.split,
                       profile_phases=args.profile_phases, kernel_flags=args.kernel_flags,
                       fused_grid=args.fused_grid)
    t0 = time.time()
    data = make_rcv1_like(args.rows, args.features, args.nnz_per_row, seed=0, shard=rank, device=dev)
    eng = DarlinEngine(cfg, comm, dev, out=sys.stderr)
    eng.load(data)
    nnz = eng.nnz
    del data
    torch.cuda.synchronize()
    t_load = time.time() - t0

    def barrier():
        comm.barrier()
        torch.cuda.synchronize()

    kkt, reset = 1e20, False
    objs, pass_wall, converged_at = [], [], None

    def one_pass(it):
        # the reference's scheduler loop (darlin.h:52-120): KKT threshold from the last violation,
        # stop when the relative objective change stays <= epsilon after a filter reset
        nonlocal kkt, reset, converged_at
        t = time.perf_counter()
        order, n_prior = eng.next_order(it)
        p = eng.run_pass(order, kkt, reset, n_prior)   # public API: H2D + pass + D2H (host sync)
        pass_wall.append(time.perf_counter() - t)
        kkt = p.violation / max(eng.num_ex_global, 1) * cfg.kkt_filter_threshold_ratio
        rel = objs[-1] / p.objective - 1 if objs else 1.0
        objs.append(p.objective)
        if 0 < rel <= args.epsilon:
            if reset and converged_at is None:
                converged_at = it
            reset = True
        else:
            reset = False
        return p

    for it in range(args.warmup):
        one_pass(it)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev_ms = 0.0
    ker_ms = 0.0
    t_wall0 = time.perf_counter()
    ev0.record()
    for it in range(args.warmup, args.warmup + args.steps):
        p = one_pass(it)
        dev_ms += p.device_ms
        ker_ms += p.kernel_ms
    ev1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    span_ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    if args.profile_phases and rank == 0:
        pr = eng.prof.cpu().tolist()
        passes = args.steps + args.warmup
        nst = max(len(eng.blocks) // eng.bps, 1)
        names = ['phase1', 'plan+prefetch', 'barrierA', 'phase3', 'barrierC']
        import collections
        mx = pr[8:8 + 2 * nst]
        p1 = [v >> 20 for v in mx[0::2]]; p3 = [v >> 20 for v in mx[1::2]]
        print('max-over-CTA phase1 us mean %.2f, phase3 %.2f (max over passes); straggler CTAs p1 %s p3 %s' % (sum(p1) / len(p1) / 1e3, sum(p3) / len(p3) / 1e3, collections.Counter(v & 0xfffff for v in mx[0::2]).most_common(4), collections.Counter(v & 0xfffff for v in mx[1::2]).most_common(4)), file=sys.stderr)
        pc = pr[8 + 2 * nst:8 + 2 * nst + 2 * eng.grid_fused]
        tot = sorted(((pc[2 * i] / passes / nst / 1e3, i, pc[2 * i + 1]) for i in range(eng.grid_fused)), reverse=True)
        print('slowest CTAs (mean phase1 us, cta, smid):', [(round(a, 2), b, c) for a, b, c in tot[:12]], 'median', round(tot[len(tot) // 2][0], 2), file=sys.stderr)
        tp = pr[8 + 2 * nst + 8192:8 + 2 * nst + 8192 + 64]
        agg = [sum(tp[8 * c + i] for c in range(8)) for i in range(8)]
        print('warp-0 grad_segment ns: z-load %.0f compute %.0f atomic %.0f | reducer: part-load %.0f update %.0f  (n=%d, reducers=%d)' % (agg[0] / max(agg[5], 1), agg[1] / max(agg[5], 1), agg[2] / max(agg[5], 1), agg[3] / max(agg[6], 1), agg[4] / max(agg[6], 1), agg[5], agg[6]), file=sys.stderr)
        sp = sorted(p1); print('phase1 max-over-CTA quantiles us:', [round(sp[int(q * (len(sp) - 1))] / 1e3, 1) for q in (0, .1, .25, .5, .75, .9, .99, 1)], file=sys.stderr)
        sp = sorted(p3); print('phase3 max-over-CTA quantiles us:', [round(sp[int(q * (len(sp) - 1))] / 1e3, 1) for q in (0, .1, .25, .5, .75, .9, .99, 1)], file=sys.stderr)
        print('per-step us (mean over CTAs):', {n: round(v / eng.grid_fused / passes / nst / 1e3, 2) for n, v in zip(names, pr)}, file=sys.stderr)

    times = torch.tensor([dev_ms, span_ms, t_wall * 1e3, ker_ms], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms, span_ms, wall_ms, ker_ms = times.tolist()
    # time to objective: keep going (untimed for the headline) until the reference's stopping rule
    it = args.warmup + args.steps

    def keep_going():
        go = converged_at is None and it < args.max_pass
        if world > 1:  # every rank sees the same objective, but a collective decision cannot diverge
            flag = torch.tensor([int(go)], dtype=torch.int64, device=dev)
            dist.broadcast(flag, 0)
            go = bool(flag.item())
        return go
    while keep_going():
        one_pass(it)
        it += 1
    n_conv = (converged_at + 1) if converged_at is not None else len(pass_wall)
    tto = torch.tensor([sum(pass_wall[:n_conv])], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tto, op=dist.ReduceOp.MAX)
    n_ex = eng.num_ex_global
    if rank == 0:
        value = n_ex * args.steps / (ker_ms / 1e3)
        e2e = n_ex * args.steps / (wall_ms / 1e3)
        out = {
            "metric": "examples/sec, L1-LR Darlin BCD (one step = one full pass over the data)",
            "value": value, "unit": "examples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ker_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp64 state / fp32 matrix values",
            "data": "synthetic rcv1-shaped (seeded), zero-init weights",
            "config": {"model": "L1 logistic regression, Darlin block coordinate descent",
                       "examples_per_gpu": args.rows, "global_examples": n_ex,
                       "features": args.features, "nnz_per_gpu": nnz,
                       "feature_blocks": len(eng.blocks), "max_block_delay": args.tau,
                       "kernel": "split (3 launches/step)" if args.split else "persistent cooperative pass kernel",
                       "lambda": 1.0, "parallelism": f"dp{world} workers x {world} key-range server shards",
                       "l2": "inputs larger than L2 (CSC index+value arrays %.0f MB per GPU)" % (nnz * 8 / 1e6),
                       "timing": "value: CUDA events around the kernels of each pass (graph replay), summed, max over ranks; e2e: wall clock incl. host scheduler, pinned H2D, D2H, sync",
                       "ms_per_step_with_copies": dev_ms / args.steps,
                       "objective_first_last": [objs[0], objs[args.warmup + args.steps - 1]],
                       "load_preprocess_s": t_load,
                       "time_to_objective": {
                           "seconds": tto.item(), "passes": n_conv, "objective": objs[n_conv - 1],
                           "converged": converged_at is not None,
                           "rule": "relative objective change <= %g on two consecutive passes (darlin.h:103-118), "
                                   "at most %d passes; wall clock of every pass from the zero model, "
                                   "incl. first-pass graph capture, H2D, D2H, host sync; max over ranks; "
                                   "data loading / preprocessing is load_preprocess_s" % (args.epsilon, args.max_pass)}},
            "clocks": clocks,
            "e2e": {"value": e2e, "unit": "examples/s", "h2d_bytes_per_step": eng.h2d_bytes_per_pass,
                    "d2h_bytes_per_step": eng.d2h_bytes_per_pass, "wall_ms_per_step": wall_ms / args.steps,
                    "api": "DarlinEngine.run_pass (pinned order+ctrl H2D, graph replay, progress D2H, host sync) per step"},
            "gpu_launches": eng.launches_per_pass * args.steps,
            "event_span_ms": span_ms,
        }
        print(json.dumps(out))
    eng.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
