"""Cost of one software grid barrier (red.release + relaxed poll) on this GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from parameter_server_b200._lib import check, cuda_lib, current_stream, ptr
from parameter_server_b200.ops._sig import apply_cuda_sigs

apply_cuda_sigs()
lib = cuda_lib()
bar = torch.zeros(4, dtype=torch.int32, device="cuda")
for grid in (74, 148, 296):
    for mode in (0, 1):
        iters = 2000
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            check(lib.psb_bench_grid_barrier(ptr(bar), iters, grid, mode, current_stream()))
            e1.record(); e1.synchronize()
        print(f"grid={grid} mode={'relaxed' if mode == 0 else 'acquire'}: {e0.elapsed_time(e1) * 1e3 / iters:.3f} us / barrier")
