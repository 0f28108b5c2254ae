"""Real multi-GPU runs (one process per GPU, torchrun): the persistent pass kernel with
cross-GPU flags over CUDA-IPC peer memory must reproduce the single-process lock-step result."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(n, script_args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29541"] + script_args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("n", [2])
def test_bench_multi_gpu_matches_emulation(n):
    if torch.cuda.device_count() < n:
        pytest.skip("not enough GPUs")
    r = _torchrun(n, ["tests/multigpu_worker.py", "--check"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res["ok"], res


@pytest.mark.parametrize("n", [2])
def test_async_sgd_multi_gpu(n):
    """async SGD over CUDA-IPC peer memory: max_delay 0 tracks the lock-step emulation (closely,
    not bitwise: owners apply pushes as they arrive), max_delay 2 (ring of 3 landing regions)
    learns; no device wait times out, no key ends up in two table slots."""
    if torch.cuda.device_count() < n:
        pytest.skip("not enough GPUs")
    r = _torchrun(n, ["tests/multigpu_sgd_worker.py"], timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert res["ok"], res
