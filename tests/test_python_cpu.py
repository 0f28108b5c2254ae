"""CPU-only checks of the Python layer: config mapping, model files, front-end helpers, generators."""
import os

import pytest

import numpy as np
import torch

from parameter_server_b200.utils import model_io
from parameter_server_b200.utils.synthetic import SparseCSR, make_ctr_like, make_rcv1_like

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_darlin_config_from_reference_conf_files():
    from parameter_server_b200.models.darlin import DarlinConfig
    c = DarlinConfig.from_proto_text(open(f"{ROOT}/example/linear/ctr/batch_l1lr.conf").read())
    assert (c.lambda_l1, c.max_block_delay, c.tail_feature_freq, c.prior_fea_group) == (10.0, 8, 4, [127, 120])
    assert c.epsilon == 2e-5 and c.countmin_n_ratio == 0.66 and c.delta_max_value == 5.0
    r = DarlinConfig.from_proto_text(open(f"{ROOT}/example/linear/rcv1/batch_l1lr.conf").read())
    assert (r.lambda_l1, r.max_pass_of_data, r.epsilon) == (1.0, 20, 1e-4)
    assert (r.feature_block_ratio, r.max_block_delay, r.random_feature_block_order) == (4.0, 0, True)  # proto defaults


def test_async_sgd_config_from_reference_conf_files():
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig
    c = AsyncSGDConfig.from_proto_text(open(f"{ROOT}/example/linear/ctr/online_l1lr.conf").read())
    assert c.algo == "FTRL" and c.minibatch == 10000 and c.num_data_pass == 10 and c.tail_feature_freq == 1
    assert (c.lambda_l1, c.lambda_l2, c.alpha, c.beta) == (10.0, 1.0, 0.01, 10.0)
    text = 'linear_method { penalty { type: L2 lambda: .5 } async_sgd { algo: STANDARD ada_grad: false } }'
    s = AsyncSGDConfig.from_proto_text(text)
    assert s.algo == "SGD" and (s.lambda_l1, s.lambda_l2) == (0.0, 0.5)


def test_model_text_roundtrip_and_file_search(tmp_path):
    keys = np.array([3, 2**63 + 5, 77, 9], dtype=np.uint64)
    vals = np.array([0.5, -1.25e-3, 0.0, np.nan])
    model_io.write_model_text(str(tmp_path / "m" / "run_S0"), keys, vals)
    model_io.write_model_text(str(tmp_path / "m" / "run_S1"), np.array([11], dtype=np.uint64), np.array([2.0]))
    text = open(tmp_path / "m" / "run_S0").read()
    assert text == "3\t0.5\n9223372036854775813\t-0.00125\n"          # zeros and NaN are skipped
    model = model_io.read_model_text(str(tmp_path / "m" / "run_S.*"))
    assert model == {3: 0.5, 2**63 + 5: -0.00125, 11: 2.0}
    assert [os.path.basename(f) for f in model_io.search_files(str(tmp_path / "m" / "run_S.*"))] == ["run_S0", "run_S1"]


def test_front_end_helpers(tmp_path):
    from parameter_server_b200.apps import ps
    for name in ("part-0", "part-1", "other"):
        (tmp_path / name).write_text("x")
    conf = f'''linear_method {{ training_data {{ format: TEXT text: SPARSE_BINARY ignore_feature_group: true
      file: "{tmp_path}/part.*" file: "{tmp_path}/other" }} }}'''
    pats = ps.conf_files(conf, "training_data")
    assert pats == [f"{tmp_path}/part.*", f"{tmp_path}/other"]
    files = ps.search_files(pats)
    assert [os.path.basename(f) for f in files] == ["other", "part-0", "part-1"]
    dc = ps.data_conf_for(conf, "training_data", files[1:])
    assert dc.startswith("format: TEXT text: SPARSE_BINARY ignore_feature_group: true") and dc.count("file:") == 2
    # two feature groups become one key space, rows keep their order
    y = torch.tensor([1.0, -1.0, 1.0])
    g1 = SparseCSR(torch.tensor([0, 1, 1, 3]), torch.tensor([10, 11, 12]), None, y, 1 << 63)
    g2 = SparseCSR(torch.tensor([0, 2, 3, 3]), torch.tensor([20, 21, 22]), torch.tensor([.5, .25, 2.0]), y, 1 << 63)
    m = ps.merge_groups({1: g1, 2: g2})
    assert m.row_ptr.tolist() == [0, 3, 4, 6]
    assert m.keys.tolist() == [10, 20, 21, 22, 11, 12] and m.vals.tolist() == [1.0, .5, .25, 2.0, 1.0, 1.0]


def test_generators_are_seeded_and_shaped():
    a = make_rcv1_like(500, 200, 10.0, seed=1, shard=2)
    b = make_rcv1_like(500, 200, 10.0, seed=1, shard=2)
    assert torch.equal(a.keys, b.keys) and torch.equal(a.vals, b.vals) and torch.equal(a.y, b.y)
    assert a.n_rows == 500 and a.row_ptr[-1] == a.nnz and int(a.keys.max()) < 200
    rows = torch.repeat_interleave(torch.arange(500), a.row_ptr[1:] - a.row_ptr[:-1])
    norms = torch.zeros(500).index_add_(0, rows, a.vals ** 2)
    assert torch.allclose(norms[norms > 0], torch.ones_like(norms[norms > 0]), atol=1e-4)   # unit-length rows
    assert set(a.y.tolist()) <= {1.0, -1.0}
    c = make_ctr_like(300, n_groups=3, keys_per_group=1 << 10, seed=2)
    assert sorted(c) == [1, 2, 3]
    for grp, d in c.items():
        assert d.vals is None and d.n_rows == 300
        assert torch.all((d.keys >> 54) == grp)          # group id in the top 10 bits (TERAFEA layout)


def test_multi_rank_protocol_model():
    """tools/protocol_check.py: random interleavings of the tagged-entry / replicated-update protocol
    of the multi-rank Darlin pass hold its invariants, and the checker notices when the protocol is broken."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("protocol_check", f"{ROOT}/tools/protocol_check.py")
    pc = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(pc)
    assert pc.check(ranks=3, keys=12, blocks=4, passes=4, schedules=60, verbose=False) == 0
    assert pc.check(ranks=2, keys=9, blocks=3, passes=5, schedules=60, seed=7, verbose=False) == 0
    assert pc.check(schedules=60, mutate="no_flow", verbose=False) > 0      # writers ignore the done flags
    assert pc.check(schedules=60, mutate="no_parity", verbose=False) > 0    # one buffer instead of two per step parity


def test_streaming_runner_ordering_model():
    """tools/runner_check.py: random interleavings of the copy lanes and the step stream of the native
    async-SGD runner (ring sets, two scratch sets, sketch order) hold the invariants with the events
    the runner records, and the checker notices each missing one."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("runner_check", f"{ROOT}/tools/runner_check.py")
    rc = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rc)
    for ring, lanes in ((2, 1), (3, 2), (4, 3), (5, 3), (5, 4), (8, 4)):
        assert rc.check(ring=ring, lanes=lanes, minibatches=30, schedules=50, verbose=False) == 0
    for lanes in (2, 3):
        assert rc.check(schedules=150, lanes=lanes, mutate="no_done", verbose=False) > 0     # ring set reused without waiting
        assert rc.check(schedules=40, lanes=lanes, mutate="no_order", verbose=False) > 0     # stage B not chained
        assert rc.check(schedules=40, lanes=lanes, mutate="one_scratch", verbose=False) > 0  # all lanes on one scratch set


def test_sgd_ctypes_mirrors_have_the_c_sizes(tmp_path):
    """The ctypes structures of models/async_sgd.py are hand-written mirrors of include/psb200.h: a
    field added on one side only shifts every later field silently.  Compare sizeof on both sides."""
    import ctypes as C
    import shutil
    import subprocess
    from parameter_server_b200.models import async_sgd as m
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "sz.cpp"
    src.write_text('#include <cstdio>\n#include "%s/include/psb200.h"\nint main() { printf("%%zu %%zu %%zu %%zu %%zu\\n", '
                   'sizeof(psb_sgd_worker), sizeof(psb_sgd_server), sizeof(psb_sgd_runner_config), '
                   'sizeof(psb_sgd_host_minibatch), sizeof(psb_sgd_ring_set)); }\n' % root)
    exe = tmp_path / "sz"
    subprocess.run(["g++", "-std=c++17", str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(t) for t in (m.SgdWorkerStruct, m.SgdServerStruct, m.RunnerConfigStruct,
                                  m.HostMinibatchStruct, m.RingSetStruct)]
    assert got == want


def test_arena_is_whole_device_pages(monkeypatch):
    """Exported allocations must be multiples of the 2 MB device page: random peer access to any other
    size was measured 30x slower (profiles/README.md section 0.5)."""
    from parameter_server_b200.parallel.symm import DEVICE_PAGE, arena_bytes
    monkeypatch.delenv("PSB_ARENA_GRANULE", raising=False)
    for n in (1, 4096, DEVICE_PAGE - 1, DEVICE_PAGE, DEVICE_PAGE + 1, (2 << 30) + 4096, 5_400_000_123):
        b = arena_bytes(n)
        assert b >= n and b % DEVICE_PAGE == 0 and b - n < DEVICE_PAGE
