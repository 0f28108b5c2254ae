"""The GPU front end on the reference-format configs: files -> native SlotReader -> device engines,
cross-checked against the CPU parameter server running the same config."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    from parameter_server_b200.utils.synthetic import make_rcv1_like, write_libsvm
    d = tmp_path_factory.mktemp("apps")
    os.makedirs(d / "train")
    for p in range(2):
        write_libsvm(make_rcv1_like(2000, 800, 20.0, seed=5, shard=p), str(d / "train" / f"part-{p}"))
    return d


def _run(conf_path):
    r = subprocess.run([sys.executable, "-m", "parameter_server_b200.apps.ps", "--app_file", str(conf_path)],
                       cwd=ROOT, capture_output=True, text=True, timeout=170)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r.stderr


def test_darlin_config_gpu_matches_cpu_cluster(dataset):
    from parameter_server_b200.runtime import host
    d = dataset
    conf = f'''linear_method {{
      training_data {{ format: TEXT text: LIBSVM file: "{d}/train/part.*" }}
      model_output {{ format: TEXT file: "{d}/model/gpu" }}
      loss {{ type: LOGIT }} penalty {{ type: L1 lambda: 0.5 }} learning_rate {{ type: CONSTANT alpha: 1 }}
      darlin {{ max_pass_of_data: 4 epsilon: 1e-9 feature_block_ratio: 0.5 random_feature_block_order: false
               local_cache {{ format: BIN file: "{d}/cache/g" }} }} }}'''
    (d / "darlin.conf").write_text(conf)
    log = _run(d / "darlin.conf")
    rows = [l.split("|") for l in log.splitlines() if l[:5].strip().isdigit()]
    gpu_obj = [float(r[1].split()[0]) for r in rows]
    cpu = host.run_cluster(conf.replace("model/gpu", "model/cpu").replace("cache/g", "cache/c"), 1, 2)
    cpu_obj = [float(l.split()[1]) for l in cpu.strip().splitlines()]
    assert len(gpu_obj) == 4 and gpu_obj == pytest.approx(cpu_obj, rel=1e-5)   # table prints 6 digits
    gm = open(d / "model" / "gpu_S0").read().split()
    cm = (open(d / "model" / "cpu_S0").read() + open(d / "model" / "cpu_S1").read()).split() \
        if os.path.exists(d / "model" / "cpu_S1") else open(d / "model" / "cpu_S0").read().split()
    assert len(gm) == len(cm) and gm[0::2] == cm[0::2]
    assert [float(x) for x in gm[1::2]] == pytest.approx([float(x) for x in cm[1::2]], rel=1e-4)


def test_async_sgd_config_and_model_evaluation(dataset):
    d = dataset
    (d / "sgd.conf").write_text(f'''linear_method {{
      training_data {{ format: TEXT text: LIBSVM file: "{d}/train/part.*" }}
      model_output {{ format: TEXT file: "{d}/model/sgd" }}
      loss {{ type: LOGIT }} penalty {{ type: L1 lambda: .01 lambda: .001 }}
      learning_rate {{ type: DECAY alpha: .5 beta: 1 }}
      async_sgd {{ algo: FTRL minibatch: 200 num_data_pass: 4 max_delay: 1 }} }}''')
    log = _run(d / "sgd.conf")
    last = [l for l in log.splitlines() if l.strip() and l.strip()[0].isdigit()][-1].split()
    assert float(last[2]) < 0.55 and float(last[3]) > 0.85
    (d / "eval.conf").write_text(f'''linear_method {{
      validation_data {{ format: TEXT text: LIBSVM file: "{d}/train/part.*" }}
      model_input {{ format: TEXT file: "{d}/model/sgd_S.*" }} }}''')
    log = _run(d / "eval.conf")
    auc = float([l for l in log.splitlines() if l.startswith("auc:")][0].split()[1])
    assert auc > 0.85


def test_shipped_ctr_batch_config_on_gpu(tmp_path):
    """example/linear/ctr/batch_l1lr.conf UNCHANGED through the GPU front end: multi-group SPARSE_BINARY
    text, prior feature groups 127 / 120 swept first without delay, tail_feature_freq 4,
    max_block_delay 8 (9 blocks per device step).  Block orders are random and drawn from different
    generators on the two sides, so the trajectories differ; both must reach the same optimum.
    (It failed on one box in round 2: the CPU half lost a whole feature group to a race in the servers'
    shared tail filter - fixed, tests/test_host_apps.py::test_tail_filter_does_not_depend_on_timing_or_worker_count.)"""
    from parameter_server_b200.runtime import host
    from parameter_server_b200.utils.synthetic import make_ctr_like, write_ps_sparse_binary
    relabel = {1: 120, 2: 127, 3: 3, 4: 5}
    os.makedirs(tmp_path / "data" / "ctr" / "train")
    for p in range(4):
        groups = make_ctr_like(1500, n_groups=4, keys_per_group=1 << 12, seed=2, shard=p, nnz_per_group=4.0)
        out = {}
        for g, d in groups.items():
            d.keys = (d.keys & ((1 << 54) - 1)) | (relabel[g] << 54)   # the group id rides in the top 10 bits
            out[relabel[g]] = d
        write_ps_sparse_binary(out, str(tmp_path / "data" / "ctr" / "train" / f"part-{p:03d}"))
    conf = open(f"{ROOT}/example/linear/ctr/batch_l1lr.conf").read()
    r = subprocess.run([sys.executable, "-m", "parameter_server_b200.apps.ps", "--app_file",
                        f"{ROOT}/example/linear/ctr/batch_l1lr.conf"], cwd=tmp_path, capture_output=True, text=True,
                       timeout=170, env=dict(os.environ, PYTHONPATH=ROOT))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    rows = [l.split("|") for l in r.stderr.splitlines() if l[:5].strip().isdigit()]
    gpu_obj = [float(x[1].split()[0]) for x in rows]
    assert "tau=8" in r.stderr and len(gpu_obj) >= 5 and gpu_obj[-1] < gpu_obj[0]
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        cpu = host.run_cluster(conf.replace("model/ctr_batch", "model/cpu").replace("cache/ctr_train", "cache/cpu"), 1, 2)
    finally:
        os.chdir(cwd)
    cpu_obj = [float(l.split()[1]) for l in cpu.strip().splitlines()]
    assert gpu_obj[-1] == pytest.approx(cpu_obj[-1], rel=2e-2), (gpu_obj, cpu_obj)
    assert os.path.exists(tmp_path / "model" / "ctr_batch_S0")
