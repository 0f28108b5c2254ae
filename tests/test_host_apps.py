"""End-to-end runs of the native parameter server (scheduler + servers + workers as threads of
one process): Darlin against the PyTorch oracle, async SGD / FTRL, model evaluation."""
import math
import os

import numpy as np
import pytest
import torch

from parameter_server_b200.runtime import host
from parameter_server_b200.utils.range import Range
from parameter_server_b200.utils.synthetic import make_rcv1_like, write_libsvm

N_FILES, ROWS, FEATS, NNZ = 4, 1500, 600, 20.0


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    d = tmp_path_factory.mktemp("rcv1")
    os.makedirs(d / "train")
    shards = []
    for p in range(N_FILES):
        s = make_rcv1_like(ROWS, FEATS, NNZ, seed=3, shard=p)
        write_libsvm(s, str(d / "train" / f"part-{p}"))
        shards.append(s)
    return d, shards


def _conf(d, body, extra_data=""):
    return f'''
linear_method {{
  training_data {{ format: TEXT text: LIBSVM file: "{d}/train/part.*" {extra_data} }}
  model_output {{ format: TEXT file: "{d}/model/m" }}
  loss {{ type: LOGIT }}
  {body}
}}'''


def _darlin_conf(d, passes=5, tau=0, random_order=False, lam=0.5, cache="c", ratio=0.5):
    return _conf(d, f'''
  penalty {{ type: L1 lambda: {lam} }}
  learning_rate {{ type: CONSTANT alpha: 1 }}
  darlin {{ max_pass_of_data: {passes} epsilon: 1e-7 feature_block_ratio: {ratio} max_block_delay: {tau}
           random_feature_block_order: {"true" if random_order else "false"}
           local_cache {{ format: BIN file: "{d}/cache/{cache}" }} }}''')


def _clear_models(d):
    """model files of an earlier run with more servers would be picked up by `m_S.*`"""
    for f in (d / "model").glob("m_*") if (d / "model").exists() else []:
        f.unlink()


def _report(text):
    rows = [l.split() for l in text.strip().splitlines()]
    return [dict(it=int(r[0]), obj=float(r[1]), rel=float(r[2]), nnz=int(r[3]), vio=float(r[4]),
                 active=int(r[5])) for r in rows]


def _oracle_trajectory(shards, n_workers, passes, lam):
    """The same schedule on the PyTorch oracle: files round robin over the workers, blocks from
    the merged slot statistics, sequential block order, KKT threshold = vio / n * 10."""
    from parameter_server_b200.models.darlin_oracle import DarlinOracle, OracleServer, run_oracle_pass
    # values are written with 6 significant digits: re-read them the way the parser sees them
    def quant(s):
        v = torch.tensor([float("%.6g" % x) for x in s.vals.tolist()], dtype=torch.float32)
        return v
    per_worker = [[] for _ in range(n_workers)]
    for i, s in enumerate(shards):
        per_worker[i % n_workers].append(s)
    dictionary = torch.unique(torch.cat([s.keys for s in shards]))
    workers = []
    for parts in per_worker:
        rows, cols, vals, ys, base = [], [], [], [], 0
        for s in parts:
            r = torch.repeat_interleave(torch.arange(s.n_rows), s.row_ptr[1:] - s.row_ptr[:-1]) + base
            rows.append(r); cols.append(torch.searchsorted(dictionary, s.keys)); vals.append(quant(s))
            ys.append(s.y); base += s.n_rows
        row, col, val, y = torch.cat(rows), torch.cat(cols), torch.cat(vals), torch.cat(ys)
        order = torch.argsort(col * base + row)
        row, col, val = row[order], col[order], val[order]
        cp = torch.zeros(dictionary.numel() + 1, dtype=torch.int64)
        cp[1:] = torch.cumsum(torch.bincount(col, minlength=dictionary.numel()), 0)
        workers.append(DarlinOracle(cp, row, val, y, base))
    nnz_ele = sum(s.nnz for s in shards)
    n_ex = sum(s.n_rows for s in shards)
    n = max(int(math.ceil(nnz_ele / n_ex * 0.5)), 1)
    kmin, kmax = int(dictionary[0]), int(dictionary[-1]) + 1
    blocks = []
    for i in range(n):
        r = Range(kmin, kmax).even_divide(n, i)
        lo = int(torch.searchsorted(dictionary, torch.tensor(r.begin)))
        hi = int(torch.searchsorted(dictionary, torch.tensor(r.end)))
        if r.end > r.begin:
            blocks.append((lo, hi))
    server = OracleServer(dictionary.numel(), "cpu", lam)
    thr, reset, out = 1e20, False, []
    for it in range(passes):
        ev = run_oracle_pass(workers, server, blocks, list(range(len(blocks))), thr, reset)
        out.append(ev)
        thr = ev["violation"] / n_ex * 10
    return out, len(blocks)


@pytest.mark.parametrize("tcp", [False, True])
def test_darlin_cluster_matches_oracle(dataset, tcp):
    d, shards = dataset
    rep = _report(host.run_cluster(_darlin_conf(d, cache="tcp" if tcp else "inproc"), 2, 2, tcp=tcp,
                                   base_port=19700 + os.getpid() % 200))
    ref, n_blocks = _oracle_trajectory(shards, 2, 5, 0.5)
    assert n_blocks == 10 and len(rep) == 5
    for a, b in zip(rep, ref):
        assert a["obj"] == pytest.approx(b["objective"], rel=1e-9)
        assert a["nnz"] == b["nnz_w"] and a["active"] == b["nnz_active"]
        assert a["vio"] == pytest.approx(b["violation"], rel=1e-9, abs=1e-12)
    assert rep[-1]["active"] < rep[0]["active"]          # the KKT filter shrank the active set
    # one model file per server, key<TAB>value, only non-zero weights
    files = sorted(os.listdir(d / "model"))
    assert files == ["m_S0", "m_S1"]
    n_lines = sum(len(open(d / "model" / f).read().splitlines()) for f in files)
    assert n_lines == rep[-1]["nnz"]


def test_darlin_bounded_delay_random_order_and_slot_cache(dataset):
    d, _ = dataset
    # the second run over the same cache prefix must hit the columnar slot cache
    host.run_cluster(_darlin_conf(d, passes=2, cache="again"), 1, 2)
    assert any(f.endswith(".info") for f in os.listdir(d / "cache"))
    base = _report(host.run_cluster(_darlin_conf(d, passes=6, cache="again", ratio=3), 2, 3))
    # 60 blocks, at most 3 in flight: staleness is small next to the number of blocks
    tau = _report(host.run_cluster(_darlin_conf(d, passes=6, tau=2, random_order=True, cache="again", ratio=3),
                                   2, 3, message_compression=False))
    objs = [r["obj"] for r in tau]
    assert all(b < a for a, b in zip(objs, objs[1:]))
    assert objs[-1] == pytest.approx(base[-1]["obj"], rel=2e-2)


def test_async_sgd_ftrl_and_model_evaluation(dataset):
    d, _ = dataset
    conf = _conf(d, '''
  penalty { type: L1 lambda: .01 lambda: .001 }
  learning_rate { type: DECAY alpha: .5 beta: 1 }
  async_sgd { algo: FTRL minibatch: 200 num_data_pass: 4 max_delay: 2 report_interval: 0 }''')
    _clear_models(d)
    loss, auc = map(float, host.run_cluster(conf, 2, 2).split())
    assert loss < 0.55 and auc > 0.85      # log 2 = 0.693 is the loss of the zero model
    ev = f'''linear_method {{
      validation_data {{ format: TEXT text: LIBSVM file: "{d}/train/part.*" }}
      model_input {{ format: TEXT file: "{d}/model/m_S.*" }} }}'''
    auc2, acc, n = host.run_cluster(ev, 0, 0).split()
    assert int(n) == N_FILES * ROWS and float(auc2) > 0.85 and float(acc) > 0.75


@pytest.mark.parametrize("algo,extra,alpha,min_auc", [
    ("STANDARD", "ada_grad: true max_delay: 1", 1, 0.88),
    ("STANDARD", "ada_grad: false max_delay: 1", .3, 0.8),
    ("FTRL", "fixing_float_by_nbytes: 2 max_delay: 1", .5, 0.88),
    ("FTRL", "fixing_float_by_nbytes: 1 tail_feature_freq: 1", .5, 0.85),   # count-min tail filter, 1-byte values
])
def test_async_sgd_variants_learn(dataset, algo, extra, alpha, min_auc):
    """Judged on the saved model over the whole training set: the loss / AUC of the last
    minibatch report is too noisy for an asynchronous run (gradients are minibatch sums)."""
    d, _ = dataset
    conf = _conf(d, f'''
  penalty {{ type: L2 lambda: .0001 }}
  learning_rate {{ type: DECAY alpha: {alpha} beta: 1 }}
  async_sgd {{ algo: {algo} {extra} minibatch: 300 num_data_pass: 3 report_interval: 0 }}''')
    _clear_models(d)
    host.run_cluster(conf, 1, 2)
    ev = f'''linear_method {{
      validation_data {{ format: TEXT text: LIBSVM file: "{d}/train/part.*" }}
      model_input {{ format: TEXT file: "{d}/model/m_S.*" }} }}'''
    auc, acc, n = host.run_cluster(ev, 0, 0).split()
    assert int(n) == N_FILES * ROWS and float(auc) > min_auc, (auc, acc)


def test_load_slot_data_roundtrip(dataset, tmp_path):
    """files -> native SlotReader (columnar cache) -> SparseCSR groups"""
    from parameter_server_b200.utils.data_io import load_slot_data
    d, shards = dataset
    conf = f'format: TEXT text: LIBSVM file: "{d}/train/part-0" file: "{d}/train/part-1"'
    groups = load_slot_data(conf, str(tmp_path / "cache_"))
    g = groups[1]
    assert g.n_rows == 2 * ROWS and g.nnz == shards[0].nnz + shards[1].nnz
    assert torch.equal(g.keys[:shards[0].nnz], shards[0].keys)
    assert torch.equal(g.y[:ROWS], shards[0].y)
    assert torch.allclose(g.vals[:shards[0].nnz], shards[0].vals, rtol=1e-5)
    # second load hits the cache and returns the same arrays
    g2 = load_slot_data(conf, str(tmp_path / "cache_"))[1]
    assert torch.equal(g2.keys, g.keys) and torch.equal(g2.row_ptr, g.row_ptr)


def test_shipped_ctr_configs_run_unchanged(tmp_path, monkeypatch):
    """example/linear/ctr/{batch,online}_l1lr.conf (SPARSE_BINARY text with feature groups, prior groups,
    tail filter, max_block_delay 8 / FTRL with ignore_feature_group) run as they are on a CPU cluster."""
    from parameter_server_b200.utils.synthetic import make_ctr_like, write_ps_sparse_binary
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    os.makedirs(tmp_path / "data" / "ctr" / "train")
    for p in range(4):
        write_ps_sparse_binary(make_ctr_like(600, seed=0, shard=p), str(tmp_path / "data" / "ctr" / "train" / f"part-{p:03d}"))
    monkeypatch.chdir(tmp_path)
    rep = _report(host.run_cluster(open(f"{root}/example/linear/ctr/batch_l1lr.conf").read(), 2, 2))
    assert len(rep) >= 5 and rep[-1]["obj"] < rep[0]["obj"] and rep[-1]["nnz"] > 0
    assert any(f.startswith("ctr_batch") or "_S" in f for f in os.listdir(tmp_path / "model"))
    loss, auc = map(float, host.run_cluster(open(f"{root}/example/linear/ctr/online_l1lr.conf").read(), 2, 2).split())
    assert math.isfinite(loss) and 0.0 <= auc <= 1.0
    os.makedirs(tmp_path / "data" / "ctr" / "test")
    write_ps_sparse_binary(make_ctr_like(500, seed=0, shard=1000), str(tmp_path / "data" / "ctr" / "test" / "part-000"))
    for name in ("eval_batch", "eval_online"):
        auc, acc, n = host.run_cluster(open(f"{root}/example/linear/ctr/{name}.conf").read(), 0, 0).split()
        assert int(n) == 500 and 0.0 <= float(auc) <= 1.0 and float(acc) > 0.5


def test_tail_filter_does_not_depend_on_timing_or_worker_count(tmp_path, monkeypatch):
    """The servers' tail filter is ONE sketch shared by all feature groups.  It used to be cleared in the
    keys round of the first group, while the count / query rounds of later groups could still be in
    flight: depending on timing the last group's query found an empty sketch and the group lost ALL its
    features (final objective 1029.26 instead of 857.14 on this data; 6 of 6 runs with two workers on
    the development machine, never on another one - which is how it surfaced, as a disagreement with
    the GPU engine).  Every run must reach the optimum the exact counts give, whatever the workers."""
    from parameter_server_b200.utils.synthetic import make_ctr_like, write_ps_sparse_binary
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    os.makedirs(tmp_path / "data" / "ctr" / "train")
    relabel = {1: 120, 2: 127, 3: 3, 4: 5}
    for p in range(4):
        out = {}
        for g, d in make_ctr_like(1500, n_groups=4, keys_per_group=1 << 12, seed=2, shard=p, nnz_per_group=4.0).items():
            d.keys = (d.keys & ((1 << 54) - 1)) | (relabel[g] << 54)
            out[relabel[g]] = d
        write_ps_sparse_binary(out, str(tmp_path / "data" / "ctr" / "train" / f"part-{p:03d}"))
    monkeypatch.chdir(tmp_path)
    conf = open(f"{root}/example/linear/ctr/batch_l1lr.conf").read()
    finals = []
    for i, workers in enumerate((1, 2, 2, 4)):
        c = conf.replace("model/ctr_batch", f"model/m{i}").replace("cache/ctr_train", f"cache/c{i}")
        finals.append(_report(host.run_cluster(c, 1, workers))[-1]["obj"])
    assert finals == pytest.approx([finals[0]] * len(finals), rel=1e-4), finals
    assert finals[0] == pytest.approx(857.144, rel=1e-3), finals   # the GPU engine (exact counts) on the same data


def test_shipped_rcv1_configs_train_then_evaluate(tmp_path, monkeypatch):
    """example/linear/rcv1: batch_l1lr.conf -> eval_batch.conf and online_l1lr.conf -> eval_online.conf,
    all four files unchanged, on synthetic data laid out the way download.sh would leave it."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for split, n, base in (("train", 4, 0), ("test", 1, 1000)):
        os.makedirs(tmp_path / "data" / "rcv1" / split)
        for p in range(n):
            write_libsvm(make_rcv1_like(1200, 400, 16.0, seed=0, shard=p + base),
                         str(tmp_path / "data" / "rcv1" / split / f"part-{p:03d}"))
    monkeypatch.chdir(tmp_path)
    conf = lambda name: open(f"{root}/example/linear/rcv1/{name}.conf").read()  # noqa: E731
    rep = _report(host.run_cluster(conf("batch_l1lr"), 2, 2))
    assert rep[-1]["obj"] < rep[0]["obj"]
    auc, acc, n = host.run_cluster(conf("eval_batch"), 0, 0).split()
    assert int(n) == 1200 and float(auc) > 0.8
    host.run_cluster(conf("online_l1lr"), 2, 2)
    auc, acc, n = host.run_cluster(conf("eval_online"), 0, 0).split()
    # the shipped config sets no max_delay: how stale the gradients of the handful of minibatches
    # are depends on thread timing (typically auc ~0.89; all of them computed at w = 0 gives ~0.65)
    assert int(n) == 1200 and float(auc) > 0.6


def test_apps_under_delivery_jitter():
    """The same app tests with PSB_VAN_JITTER_US: every delivery of the in-process van waits a random
    time first, so the nodes interleave differently from run to run while each inbox keeps its order.
    An ordering assumption between the rounds of a protocol (such as the tail-filter race) shows up
    here instead of on somebody else's machine.  Runs in a subprocess: the van reads the variable once."""
    import subprocess
    import sys
    if os.environ.get("PSB_VAN_JITTER_US"):
        pytest.skip("already running under jitter")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-p", "no:cacheprovider",
                        "-k", "oracle or tail_filter or bounded_delay or ftrl_and_model_evaluation"],
                       env=dict(os.environ, PSB_VAN_JITTER_US="400"), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
    assert " passed" in r.stdout and "failed" not in r.stdout
