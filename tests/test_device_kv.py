"""The public KV API (include/ps.h: KVVector / KVBufferedVector / KVMap, push / pull with
timestamps and wait_time, operators, received(t)) on both data planes.

build/kvops and build/hello are written against ps.h only.  With `-device gpu` the same binaries
keep their shards in HBM, match / apply / gather with the kernels of devkv.cu and move value
frames through CUDA-IPC peer memory (PeerVan); the results must be those of the CPU run.
ref: src/app/hello_world/main.cc, src/parameter/shared_parameter.h:17-40, kv_map.h:4-45
"""
import os
import random
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module", autouse=True)
def _binaries():
    if not (ROOT / "build" / "kvops").exists():
        from parameter_server_b200 import build
        build.build_host()


def _run(binary, servers, workers, *flags, timeout=180):
    env = dict(os.environ, PS_BASE_PORT=str(random.randint(10000, 30000)))
    r = subprocess.run([str(ROOT / "script" / "local.sh"), str(ROOT / "build" / binary), str(servers),
                        str(workers), *flags], cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout, r.stderr


def _kvops_lines(out):
    return sorted(l for l in out.splitlines() if l[:1] in "SW" and ":" in l and ("ok" in l or "buffered" in l or "WRONG" in l))


def test_kvops_host():
    out, _ = _run("kvops", 2, 2)
    lines = _kvops_lines(out)
    assert lines == ["S0: buffered 751.0", "S1: buffered 749.0", "W0: vec ok map ok", "W1: vec ok map ok"], out


@pytest.mark.gpu
def test_kvops_device_matches_host():
    host, _ = _run("kvops", 3, 2)
    dev, err = _run("kvops", 3, 2, "-device", "gpu")
    assert _kvops_lines(dev) == _kvops_lines(host), dev + err
    assert "(HBM)" in dev + err and "WRONG" not in dev


@pytest.mark.gpu
def test_hello_world_with_shards_in_hbm():
    """bin/hello.cc unchanged (the reference's hello_world on ps.h), KVVector storage in HBM."""
    out, _ = _run("hello", 2, 2, "-device", "gpu")
    got = {l.split(":")[0]: " ".join(l.split(":", 1)[1].split()) for l in out.splitlines() if "pulled weight" in l}
    assert got == {
        "W0": "pulled weight [4]: 0 0.2 0.4 0.5 at key [4]: 0 2 4 5",
        "W1": "pulled weight [4]: 0 0.1 0.3 0.4 at key [4]: 0 1 3 4",
    }, out


@pytest.mark.gpu
def test_traffic_goes_through_peer_memory():
    """With -traffic_statistics the van reports its bytes; in gpu mode the staged frames are counted too."""
    _, err = _run("kvops", 2, 2, "-device", "gpu", "-traffic_statistics")
    assert "Gbyte" in err


def _darlin_table(err):
    """rows of the per-pass table the scheduler prints: (iter, objective, |w|_0, #active)"""
    rows = []
    for l in err.splitlines():
        p = l.split("|")
        if len(p) >= 4 and p[0].strip().isdigit():
            rows.append((int(p[0]), p[1].split()[0], int(p[2]), int(p[3].split()[1])))
    return rows


@pytest.mark.gpu
def test_darlin_binary_with_model_shards_in_hbm(tmp_path):
    """build/ps (the C++ scheduler / servers / workers) on a Darlin config: with -device gpu the
    model shards are HBM columns and (G, U) / weight blocks travel through peer memory; the
    objective table must be the one of the CPU run."""
    from parameter_server_b200.utils.synthetic import make_rcv1_like, write_libsvm
    os.makedirs(tmp_path / "train")
    for p in range(4):
        write_libsvm(make_rcv1_like(1500, 600, 20.0, seed=3, shard=p), str(tmp_path / "train" / f"part-{p}"))
    tables = {}
    for mode in ("cpu", "gpu"):
        conf = tmp_path / f"{mode}.conf"
        conf.write_text(f'''
linear_method {{
  training_data {{ format: TEXT text: LIBSVM file: "{tmp_path}/train/part.*" }}
  model_output {{ format: TEXT file: "{tmp_path}/model_{mode}/m" }}
  loss {{ type: LOGIT }}
  penalty {{ type: L1 lambda: 0.5 }}
  learning_rate {{ type: CONSTANT alpha: 1 }}
  darlin {{ max_pass_of_data: 4 epsilon: 1e-7 feature_block_ratio: 0.5 max_block_delay: 0
           random_feature_block_order: false local_cache {{ format: BIN file: "{tmp_path}/cache_{mode}/c" }} }}
}}''')
        _, err = _run("ps", 2, 2, "-app_file", str(conf), "-device", mode, timeout=300)
        tables[mode] = _darlin_table(err)
        assert len(tables[mode]) == 4, err[-2000:]
    assert tables["gpu"] == tables["cpu"]
    assert float(tables["gpu"][-1][1]) < float(tables["gpu"][0][1])
