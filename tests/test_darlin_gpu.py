"""Darlin device kernels vs the plain-PyTorch fp64 restatement of the reference loops."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _oracle_from_csr(data, dictionary, device):
    """Plain CSC (no label folding) of one worker in the global column space, via torch only."""
    from parameter_server_b200.models.darlin_oracle import DarlinOracle
    keys = data.keys.to(device)
    col = torch.searchsorted(dictionary, keys)
    keep = (col < dictionary.numel())
    keep &= dictionary[col.clamp(max=dictionary.numel() - 1)] == keys
    n_rows = data.n_rows
    row = torch.repeat_interleave(torch.arange(n_rows, device=device),
                                  (data.row_ptr[1:] - data.row_ptr[:-1]).to(device))
    row, col = row[keep], col[keep]
    val = None if data.vals is None else data.vals.to(device)[keep]
    order = torch.argsort(col * n_rows + row)
    row, col = row[order], col[order]
    val = None if val is None else val[order]
    n_cols = dictionary.numel()
    col_ptr = torch.zeros(n_cols + 1, dtype=torch.int64, device=device)
    col_ptr[1:] = torch.cumsum(torch.bincount(col, minlength=n_cols), 0)
    return DarlinOracle(col_ptr, row, val, data.y.to(device), n_rows)


def _run_compare(world, binary, tau, passes=4, n_rows=3000, n_feat=700, lam=0.5, tail=0, seed=3,
                 fused=False, graph=False, expect_filter=True, emulated=False, repeat_blocks=False,
                 grid_per_rank=0):
    from parameter_server_b200.models.darlin import (DarlinConfig, DarlinEngine, run_pass_emulated,
                                                     run_pass_lockstep)
    from parameter_server_b200.models.darlin_oracle import OracleServer, run_oracle_pass
    from parameter_server_b200.parallel.comm import run_threads
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    cfg = DarlinConfig(lambda_l1=lam, max_block_delay=tau, tail_feature_freq=tail,
                       feature_block_ratio=0.5, seg_len=64, use_graph=graph, seed=seed,
                       fused=fused or emulated)
    shards = [make_rcv1_like(n_rows, n_feat, 24.0, seed=seed, shard=r, device=dev, binary=binary)
              for r in range(world)]

    def build(comm):
        e = DarlinEngine(cfg, comm, dev)
        e.load(shards[comm.rank])
        return e

    engines = run_threads(world, build)
    e0 = engines[0]
    dictionary = e0.dicts[1]
    # sanity on the partition
    assert sum(e0.shard_len) == dictionary.numel()
    workers = [_oracle_from_csr(shards[r], dictionary, dev) for r in range(world)]
    for wk in workers:
        wk.delta_max = cfg.delta_max_value
    server = OracleServer(dictionary.numel(), dev, lam, cfg.eta, cfg.delta_init_value,
                          cfg.delta_max_value)
    blocks = [(b[2], b[3]) for b in e0.blocks]
    assert len(blocks) >= 3
    thr, reset = 1e20, False
    rng = np.random.RandomState(1)
    n_ex = sum(s.n_rows for s in shards)
    for it in range(passes):
        order = list(range(len(blocks)))
        rng.shuffle(order)
        if repeat_blocks:  # the same block in consecutive steps and two steps apart (a, a, b, a, ...)
            order = [order[0], order[0], order[1], order[0]] + order
        if emulated:
            p = run_pass_emulated(engines, order, thr, reset, grid_per_rank=grid_per_rank)
        elif fused:
            assert world == 1
            p = e0.run_pass(order, thr, reset)
        else:
            p = run_pass_lockstep(engines, order, thr, reset)
        ref = run_oracle_pass(workers, server, blocks, order, thr, reset, e0.bps)
        assert p.nnz_w == ref["nnz_w"], (it, p.nnz_w, ref["nnz_w"])
        assert p.nnz_active_set == ref["nnz_active"], (it, p.nnz_active_set, ref["nnz_active"])
        assert abs(p.objective - ref["objective"]) <= 1e-9 * abs(ref["objective"]), (it, p.objective, ref["objective"])
        assert abs(p.violation - ref["violation"]) <= 1e-9 * max(1.0, ref["violation"])
        # aggressive threshold so that the KKT filter really triggers in the test
        thr = p.violation / n_ex * (0.5 if it < 2 else 10.0)
        reset = it == 2
    # weights, shard by shard
    w_dev = torch.cat([e.shard_weights() for e in engines])
    both_nan = torch.isnan(w_dev) & torch.isnan(server.w)
    assert torch.equal(torch.isnan(w_dev), torch.isnan(server.w))
    assert torch.allclose(torch.where(both_nan, 0, w_dev), torch.where(both_nan, 0, server.w),
                          rtol=1e-9, atol=1e-12)
    if expect_filter:
        assert int(torch.isnan(server.w).sum()) > 0, "test should exercise the KKT filter"
    if emulated:  # every rank's replica is bit-identical to the shards
        for e in engines:
            rep = e.w_cache[:dictionary.numel()]
            assert torch.equal(torch.isnan(rep), torch.isnan(w_dev))
            assert torch.equal(torch.where(torch.isnan(rep), 0, rep), torch.where(both_nan, 0, w_dev))
    for e in engines:
        e.close()


def test_darlin_single_rank_matches_oracle():
    _run_compare(world=1, binary=False, tau=0)


def test_darlin_binary_features():
    _run_compare(world=1, binary=True, tau=0, lam=2.0)


def test_darlin_bounded_delay_steps():
    _run_compare(world=1, binary=False, tau=2)


def test_darlin_two_emulated_ranks():
    _run_compare(world=2, binary=False, tau=0)


def test_darlin_three_emulated_ranks_delay_tail():
    _run_compare(world=3, binary=False, tau=1, tail=2)


def test_darlin_fused_persistent_kernel_matches_oracle():
    _run_compare(world=1, binary=False, tau=0, fused=True)


def test_darlin_fused_binary_delay_graph():
    _run_compare(world=1, binary=True, tau=3, lam=2.0, fused=True, graph=True)


def test_darlin_fused_large_grid_many_segments():
    # more segments than resident warps and long multi-segment columns
    _run_compare(world=1, binary=False, tau=1, n_rows=60000, n_feat=300, fused=True,
                 expect_filter=False)


# The persistent multi-rank pass (tagged entries, replicated update, flow-control flags): all
# ranks run side by side in ONE cooperative launch, the code path of the multi-GPU run.
def test_darlin_multirank_pass_two_ranks():
    _run_compare(world=2, binary=False, tau=0, emulated=True)


def test_darlin_multirank_pass_three_ranks_delay_tail():
    _run_compare(world=3, binary=False, tau=1, tail=2, emulated=True)


def test_darlin_multirank_pass_four_ranks_binary_repeated_blocks():
    _run_compare(world=4, binary=True, tau=0, lam=2.0, emulated=True, repeat_blocks=True)


def test_darlin_multirank_pass_many_segments_small_grid():
    # more segments than resident warps: phase-1 state parked in the stash
    _run_compare(world=2, binary=False, tau=1, n_rows=40000, n_feat=300, emulated=True,
                 expect_filter=False, grid_per_rank=6)


def test_darlin_graph_train_decreases_objective(tmp_path):
    from parameter_server_b200.models.darlin import DarlinConfig, DarlinEngine
    from parameter_server_b200.utils.model_io import read_model_text
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    data = make_rcv1_like(20000, 3000, 40.0, seed=5, device=dev)
    cfg = DarlinConfig(lambda_l1=1.0, max_pass_of_data=6, use_graph=True)
    e = DarlinEngine(cfg, device=dev, out=open(tmp_path / "log.txt", "w"))
    e.load(data)
    prog = e.train()
    objs = [p.objective for p in prog]
    assert all(b < a for a, b in zip(objs, objs[1:])), objs
    assert objs[0] < 20000 * np.log(2)
    path = e.save_model(str(tmp_path / "model"))
    m = read_model_text([path])
    assert len(m) == prog[-1].nnz_w
    e.close()


def test_darlin_countmin_tail_filter_is_superset_of_exact():
    """tail_filter="countmin" (the reference's sketch) never drops a feature the exact count keeps."""
    from parameter_server_b200.models.darlin import DarlinConfig, DarlinEngine
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    data = make_rcv1_like(3000, 4000, 12.0, seed=9, device=dev)
    dicts = {}
    for mode in ("exact", "countmin"):
        eng = DarlinEngine(DarlinConfig(tail_feature_freq=3, tail_filter=mode, countmin_k=2), None, dev)
        eng.load(data)
        dicts[mode] = set(eng.keys_all.tolist())
        eng.close()
    assert dicts["exact"] <= dicts["countmin"]
    extra = len(dicts["countmin"]) - len(dicts["exact"])
    # 2 probes into 2x as many cells as keys: collisions keep roughly a tenth of the rare features
    assert 0 < len(dicts["exact"]) < 4000 and extra <= 0.25 * len(dicts["exact"])


def test_darlin_multirank_fp32_push_stays_close_and_consistent():
    """push_format="fp32": the pushed (g, u) are rounded to fp32 (the lossy push filter); the
    trajectory stays within fp32 rounding of the exact one and all replicas remain bit-identical."""
    from parameter_server_b200.models.darlin import DarlinConfig, DarlinEngine, run_pass_emulated
    from parameter_server_b200.parallel.comm import run_threads
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    shards = [make_rcv1_like(3000, 700, 24.0, seed=3, shard=r, device=dev) for r in range(2)]
    objs = {}
    for fmt in ("fp64", "fp32"):
        cfg = DarlinConfig(lambda_l1=0.5, feature_block_ratio=0.5, seg_len=64, use_graph=False, seed=3,
                           push_format=fmt)

        def build(comm):
            e = DarlinEngine(cfg, comm, dev)
            e.load(shards[comm.rank])
            return e
        engines = run_threads(2, build)
        thr, out = 1e20, []
        for it in range(4):
            p = run_pass_emulated(engines, list(range(len(engines[0].blocks))), thr, False)
            out.append(p.objective)
            thr = p.violation / 6000 * 10
        assert torch.equal(engines[0].w_cache.nan_to_num(nan=-7.0), engines[1].w_cache.nan_to_num(nan=-7.0))
        objs[fmt] = out
        for e in engines:
            e.close()
    assert objs["fp32"] != objs["fp64"]
    for a, b in zip(objs["fp32"], objs["fp64"]):
        assert abs(a - b) <= 1e-5 * abs(b)
