"""Dense feature-block tile (tcgen05 + TMA + TMEM, csrc/cuda/dense_tile.cu) against its roofline.

    python benchmarks/dense_tile_bench.py [--n 1048576] [--d 1024] [--k 64]

Scoring K models on n examples with d dense features reads X once: n*d*2 bytes (bf16) dominate, so
the bound is HBM bandwidth (MEASURED_PEAKS.json hbm_gbs), not the tensor pipe; cuBLAS (torch.matmul
on the same operands, no loss epilogue) is printed next to it.  CUDA events, 3 warm-ups, inputs
larger than L2 (2 GiB at the defaults), median of --iters runs.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def timed(fn, iters):
    ts = []
    for i in range(iters + 3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 20)
    ap.add_argument("--d", type=int, default=1024)
    ap.add_argument("--k", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    from parameter_server_b200.ops.dense import dense_margins
    dev = torch.device("cuda:0")
    x = (torch.randn(a.n, a.d, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn(a.k, a.d, device=dev) * 0.1).to(torch.bfloat16)
    y = torch.where(torch.rand(a.n, device=dev) > 0.5, 1.0, -1.0)
    ms = timed(lambda: dense_margins(x, w, y), a.iters)
    ms_loss = timed(lambda: dense_margins(x, w, y, with_margins=False), a.iters)
    ms_lib = timed(lambda: torch.matmul(x, w.T), a.iters)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    bytes_moved = a.n * a.d * 2 + a.k * a.d * 2 + a.n * a.k * 4 + a.n * 4
    out, loss = dense_margins(x, w, y)
    ref = x.float() @ w.float().T
    print(json.dumps({"n": a.n, "d": a.d, "k": a.k, "ms": ms, "ms_loss_only": ms_loss, "cublas_ms_no_epilogue": ms_lib,
                      "loss_only_frac_of_measured_hbm": (a.n * a.d * 2 + a.n * 4) / ms_loss / 1e6 / hbm,
                      "gbs": bytes_moved / ms / 1e6, "frac_of_measured_hbm": bytes_moved / ms / 1e6 / hbm,
                      "tflops": 2.0 * a.n * a.d * a.k / ms / 1e9,
                      "max_abs_err_vs_fp32": float((out - ref).abs().max())}))


if __name__ == "__main__":
    main()
