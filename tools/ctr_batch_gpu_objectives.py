"""The GPU half of tests/test_apps_gpu.py::test_shipped_ctr_batch_config_on_gpu on its own: generate the
test's data, run example/linear/ctr/batch_l1lr.conf through the GPU front end, print the objective of
every pass.  The test compares the last value with the CPU cluster's to 2 %.  Measured on a one-GPU
box: 857.144 after 11 passes, 4 433 features after the tail filter (the exact count of keys seen more
than 4 times) - the value the CPU cluster reaches too since the race in its shared tail filter was
fixed.

    python tools/ctr_batch_gpu_objectives.py
"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    from parameter_server_b200.utils.synthetic import make_ctr_like, write_ps_sparse_binary
    tmp = tempfile.mkdtemp(prefix="ctr_batch_")
    os.makedirs(os.path.join(tmp, "data", "ctr", "train"))
    relabel = {1: 120, 2: 127, 3: 3, 4: 5}
    for p in range(4):
        groups = make_ctr_like(1500, n_groups=4, keys_per_group=1 << 12, seed=2, shard=p, nnz_per_group=4.0)
        out = {}
        for g, d in groups.items():
            d.keys = (d.keys & ((1 << 54) - 1)) | (relabel[g] << 54)
            out[relabel[g]] = d
        write_ps_sparse_binary(out, os.path.join(tmp, "data", "ctr", "train", f"part-{p:03d}"))
    r = subprocess.run([sys.executable, "-m", "parameter_server_b200.apps.ps", "--app_file",
                        os.path.join(ROOT, "example", "linear", "ctr", "batch_l1lr.conf")], cwd=tmp,
                       capture_output=True, text=True, timeout=170, env=dict(os.environ, PYTHONPATH=ROOT))
    rows = [l for l in r.stderr.splitlines() if l[:5].strip().isdigit()]
    import torch
    print("visible GPUs:", torch.cuda.device_count(), "cpus:", os.cpu_count(), "rc:", r.returncode)
    for l in rows:
        print(l)
    if r.returncode:
        print(r.stderr[-2000:])
    return r.returncode


if __name__ == "__main__":
    sys.exit(main())
