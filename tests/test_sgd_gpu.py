"""Async SGD device path (sgd.cu) and the evaluation / sketch / codec kernels (eval.cu)."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_auc_accuracy_logloss_match_references():
    from sklearn.metrics import roc_auc_score
    from parameter_server_b200.ops import evaluation as E
    g = torch.Generator(device="cuda").manual_seed(0)
    y = torch.where(torch.rand(50000, generator=g, device="cuda") > 0.3, 1.0, -1.0)
    p = y * 0.7 + torch.randn(50000, generator=g, device="cuda")
    a = E.auc(y, p)
    assert a == pytest.approx(roc_auc_score((y > 0).cpu().numpy(), p.cpu().numpy()), abs=1e-6)
    acc, ll = E.accuracy_logloss(y, p)
    assert acc == pytest.approx(float(((p > 0) == (y > 0)).float().mean()), abs=1e-7)
    assert ll == pytest.approx(float(torch.log1p(torch.exp(-(y * p).double())).mean()), rel=1e-9)


def test_countmin_device_equals_host_sketch():
    from parameter_server_b200.ops import evaluation as E
    from parameter_server_b200.runtime import host
    rng = np.random.RandomState(0)
    keys = np.unique(rng.randint(0, 2**62, 20000).astype(np.uint64))
    counts = rng.randint(1, 9, keys.size).astype(np.uint8)
    keep_h, table_h = host.countmin_filter(keys, counts, n=40000, k=2, freq=4)
    cm = E.CountMin(40000, 2, "cuda")
    kd = torch.from_numpy(keys.view(np.int64)).cuda()
    cm.insert(kd, torch.from_numpy(counts.astype(np.int32)).cuda())
    assert np.array_equal(cm.table.cpu().numpy(), table_h[:cm.n])
    assert np.array_equal(cm.query_gt(kd, 4).cpu().numpy(), keep_h)


@pytest.mark.parametrize("nbytes", [1, 2])
def test_fixfloat_device_codec(nbytes):
    from parameter_server_b200.ops import evaluation as E
    g = torch.Generator(device="cuda").manual_seed(1)
    x = (torch.rand(100000, generator=g, device="cuda") * 180 - 90).float()
    code = E.fixfloat_encode(x, nbytes, -90, 90, seed=5)
    y = E.fixfloat_decode(code, nbytes, -90, 90)
    step = 180.0 / (2 ** (8 * nbytes) - 2)
    assert float((y - x).abs().max()) <= step * 1.001 + 1e-4
    assert abs(float((y - x).double().mean())) < step * 0.02
    # the byte stream is the wire format of the host filter: decode with its formula
    c = code.cpu().numpy().reshape(-1, nbytes).astype(np.uint64)
    r = sum(c[:, j] << np.uint64(8 * j) for j in range(nbytes)).astype(np.float64)
    assert np.allclose(r / (2 ** (8 * nbytes) - 2) * 180 - 90, y.cpu().numpy(), atol=1e-4)


def test_fill_and_popcount():
    from parameter_server_b200.ops import evaluation as E
    t = torch.empty(200000, dtype=torch.float64, device="cuda")
    E.fill_(t, "GAUSSIAN", mean=2.0, std=3.0, seed=9)
    assert float(t.mean()) == pytest.approx(2.0, abs=0.05) and float(t.std()) == pytest.approx(3.0, abs=0.05)
    assert float(E.fill_(t, "CONSTANT", constant=0.25).min()) == 0.25 and float(E.fill_(t).abs().max()) == 0
    w = torch.randint(-2**31, 2**31 - 1, (10000,), device="cuda", dtype=torch.int32)
    ref = sum(bin(v & 0xFFFFFFFF).count("1") for v in w.cpu().tolist())
    assert E.bitmap_popcount(w) == ref


def _reference_sgd(shards, cfg, steps, world):
    """Sequential float32 restatement: per step every worker computes its gradient on the same
    weights, the server applies the pushes in worker order (FTRL / AdaGrad / SGD)."""
    state = {}
    losses = []

    def prox(z, eta):
        l = cfg.lambda_l1 * eta
        if -l <= z <= l:
            return 0.0
        return (z - l) / (1 + cfg.lambda_l2 * eta) if z > 0 else (z + l) / (1 + cfg.lambda_l2 * eta)

    f = np.float32
    const_lr = cfg.lr_type == "CONSTANT"  # learning_rate.h:15-22: eta = alpha
    for step in range(steps):
        pushes = []
        tot_loss, tot_rows = 0.0, 0
        for s in shards:
            a = min(step * cfg.minibatch, s.n_rows)
            b = min(a + cfg.minibatch, s.n_rows)
            lo, hi = int(s.row_ptr[a]), int(s.row_ptr[b])
            keys = s.keys[lo:hi].cpu().numpy()
            vals = s.vals[lo:hi].cpu().numpy()
            rows = np.repeat(np.arange(b - a), (s.row_ptr[a + 1:b + 1] - s.row_ptr[a:b]).cpu().numpy())
            y = s.y[a:b].cpu().numpy()
            w = np.array([state.get(int(k), (0, 0, 0))[0] for k in keys], dtype=f)
            m = np.zeros(b - a, dtype=np.float64)
            np.add.at(m, rows, (w * vals).astype(np.float64))
            m = m.astype(f)
            ym = y * m
            tot_loss += float(np.log1p(np.exp(-ym.astype(np.float64))).sum())
            tot_rows += b - a
            res = (-y / (1 + np.exp(ym))).astype(f)
            uniq, inv = np.unique(keys, return_inverse=True)
            g = np.zeros(uniq.size, dtype=np.float64)
            np.add.at(g, inv, (res[rows] * vals).astype(np.float64))
            pushes.append((uniq, g.astype(f)))
        losses.append(tot_loss / max(tot_rows, 1))
        for uniq, g in pushes:
            for k, gi in zip(uniq.tolist(), g.tolist()):
                w, z, n = state.get(k, (f(0), f(0), f(0)))
                gi = f(gi)
                if cfg.algo == "FTRL":
                    n2 = f(math.sqrt(n * n + gi * gi))
                    z = f(z + gi - (n2 - n) / f(cfg.alpha) * w)
                    n = n2
                    eta = f(cfg.alpha) if const_lr else f(cfg.alpha) / (n + f(cfg.beta))
                    w = f(prox(-z * eta, eta))
                elif cfg.algo == "ADAGRAD":
                    z = f(z + gi * gi)
                    eta = f(cfg.alpha) if const_lr else f(cfg.alpha) / (f(math.sqrt(z)) + f(cfg.beta))
                    w = f(prox(w - eta * gi, eta))
                else:
                    eta = f(cfg.alpha) if const_lr else f(cfg.alpha) / (f(math.sqrt(step + 1)) + f(cfg.beta))
                    w = f(prox(w - eta * gi, eta))
                state[k] = (w, z, n)
    return losses, {k: v[0] for k, v in state.items() if v[0] != 0}


def _run_engines(world, algo, quant=0, tau=0, steps=12, lr_type="DECAY", alpha=0.5, push_format="int8"):
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
    from parameter_server_b200.parallel.comm import run_threads
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    cfg = AsyncSGDConfig(algo=algo, minibatch=256, alpha=alpha, beta=1.0, lambda_l1=0.02, lambda_l2=0.01,
                         max_delay=tau, fixing_float_by_nbytes=quant, table_capacity=1 << 14, lr_type=lr_type,
                         push_format=push_format)
    shards = [make_rcv1_like(steps * 256, 3000, 24.0, seed=4, shard=r, device=dev) for r in range(world)]
    # spread the keys over the whole uint64 range so that every shard owns some
    for s in shards:
        s.keys = (s.keys * ((1 << 63) // 3000 * 2 // 1)).contiguous()

    def build(comm):
        e = AsyncSGDEngine(cfg, comm, dev)
        e.load(shards[comm.rank])
        return e

    engines = run_threads(world, build)
    losses = []
    for step in range(steps):
        for e in engines:
            e.worker_step(step)
        for e in engines:
            e.server_step()
        torch.cuda.synchronize()
        tot = sum(e.wstats.cpu().tolist()[0] for e in engines)
        rows = sum(e.wstats.cpu().tolist()[2] for e in engines)
        for e in engines:
            e.wstats.zero_()
        losses.append(tot / rows)
    model = {}
    for e in engines:
        assert int(e.overflow.item()) == 0
        k, w = e.shard_model()
        model.update(dict(zip((x & 0xFFFFFFFFFFFFFFFF for x in k.cpu().tolist()), w.cpu().tolist())))
    for e in engines:
        e.close()
    return cfg, shards, losses, model


@pytest.mark.parametrize("world,algo", [(1, "FTRL"), (1, "ADAGRAD"), (1, "SGD"), (3, "FTRL")])
def test_sgd_engine_matches_sequential_reference(world, algo):
    cfg, shards, losses, model = _run_engines(world, algo)
    for s in shards:
        s.keys = s.keys  # already scaled
    ref_losses, ref_model = _reference_sgd(shards, cfg, len(losses), world)
    assert np.allclose(losses, ref_losses, rtol=2e-4, atol=1e-6)
    assert losses[-1] < losses[0]
    keys = sorted(ref_model)
    ref_model = {k & 0xFFFFFFFFFFFFFFFF: v for k, v in ref_model.items()}
    assert set(model) == set(ref_model)
    a = np.array([model[k] for k in sorted(model)])
    b = np.array([ref_model[k] for k in sorted(model)])
    assert np.allclose(a, b, rtol=5e-3, atol=2e-5)


def test_sgd_int8_push_and_bounded_delay_ring():
    _, _, exact, _ = _run_engines(2, "FTRL", quant=0, tau=0)
    _, _, q8, _ = _run_engines(2, "FTRL", quant=1, tau=2)
    # block-scaled int8 gradients track the fp32 run closely and still learn
    assert q8[-1] < q8[0] and abs(q8[-1] - exact[-1]) < 0.03


@pytest.mark.parametrize("algo", ["FTRL", "ADAGRAD", "SGD"])
def test_sgd_constant_learning_rate(algo):
    """learning_rate { type: CONSTANT }: eta = alpha in all three server updates (learning_rate.h:15-22)"""
    cfg, shards, losses, model = _run_engines(1, algo, lr_type="CONSTANT", alpha=0.05)
    ref_losses, ref_model = _reference_sgd(shards, cfg, len(losses), 1)
    assert np.allclose(losses, ref_losses, rtol=2e-4, atol=1e-6)
    ref_model = {k & 0xFFFFFFFFFFFFFFFF: v for k, v in ref_model.items()}
    assert set(model) == set(ref_model)
    a = np.array([model[k] for k in sorted(model)])
    b = np.array([ref_model[k] for k in sorted(model)])
    assert np.allclose(a, b, rtol=5e-3, atol=2e-5)
    # and it is a different trajectory from DECAY with the same alpha
    _, _, decay, _ = _run_engines(1, algo, lr_type="DECAY", alpha=0.05)
    assert not np.allclose(losses[2:], decay[2:], rtol=1e-3)


def test_sgd_fp8_push_tracks_fp32():
    _, _, exact, _ = _run_engines(2, "FTRL", quant=0, tau=0)
    _, _, q8, _ = _run_engines(2, "FTRL", quant=1, tau=0, push_format="fp8")
    assert q8[-1] < q8[0] and abs(q8[-1] - exact[-1]) < 0.03


def test_sgd_tail_filter_is_the_persistent_host_sketch():
    """tail_feature_freq: keys are kept once their CUMULATIVE count over the minibatches seen so far
    exceeds the bound (one CountMin per worker, src/learner/sgd.h:137-147) - checked against the
    host sketch fed with the same sequence."""
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
    from parameter_server_b200.runtime import host
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    cfg = AsyncSGDConfig(minibatch=200, tail_feature_freq=2, countmin_n=50000, countmin_k=2, table_capacity=1 << 14)
    data = make_rcv1_like(1000, 2000, 10.0, seed=8, device=dev)
    eng = AsyncSGDEngine(cfg, None, dev)
    eng.load(data)
    seen_k, seen_c = [], []
    kept_any_late = False
    for mb in range(5):
        m = eng._prepare(mb)
        a, b = mb * 200, (mb + 1) * 200
        lo, hi = int(data.row_ptr[a]), int(data.row_ptr[b])
        uniq, cnt = np.unique(data.keys[lo:hi].cpu().numpy().view(np.uint64), return_counts=True)
        seen_k.append(uniq); seen_c.append(np.minimum(cnt, 255).astype(np.uint8))
        keep, _ = host.countmin_filter(np.concatenate(seen_k), np.concatenate(seen_c), n=eng.tail_sketch.n, k=2, freq=2)
        want = uniq[keep[-uniq.size:]]
        assert np.array_equal(m.kept_keys().cpu().numpy().view(np.uint64), want), mb
        # a key that shows up once per minibatch is dropped at first and kept later
        kept_any_late |= mb >= 2 and bool(np.any(cnt[keep[-uniq.size:]] == 1))
    assert kept_any_late
    eng.close()


def test_sgd_streaming_equals_resident():
    """train_stream (pinned host minibatches, prefetch thread, H2D + localisation on a side stream)
    walks the same trajectory as the resident engine on the same minibatches."""
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
    from parameter_server_b200.utils.synthetic import SparseCSR, make_rcv1_like
    dev = torch.device("cuda:0")
    data = make_rcv1_like(8 * 256, 3000, 24.0, seed=4, device="cpu")
    cfg = AsyncSGDConfig(algo="FTRL", minibatch=256, alpha=0.5, lambda_l1=0.02, table_capacity=1 << 14)

    def host_batches():
        for a in range(0, data.n_rows, 256):
            lo, hi = int(data.row_ptr[a]), int(data.row_ptr[a + 256])
            yield SparseCSR((data.row_ptr[a:a + 257] - lo).pin_memory(), data.keys[lo:hi].pin_memory(),
                            data.vals[lo:hi].pin_memory(), data.y[a:a + 256].pin_memory(), data.n_features)

    res = AsyncSGDEngine(cfg, None, dev)
    res.load(data)
    for i in range(8):
        res.step(i)
    k0, w0 = res.shard_model()
    loss0 = res.wstats.cpu().tolist()[0]
    res.close()
    st = AsyncSGDEngine(cfg, None, dev)
    st.open(256 * 64)
    assert st.train_stream(host_batches(), 10) == 8 * 256   # two steps past the end push nothing
    k1, w1 = st.shard_model()
    assert st.h2d_bytes > 0 and torch.equal(k0, k1) and torch.allclose(w0, w1, rtol=1e-6, atol=1e-7)
    assert st.wstats.cpu().tolist()[0] == pytest.approx(loss0, rel=1e-6)
    st.close()


def test_sgd_localize_matches_numpy():
    """psb_sgd_localize (one enqueue, all sizes on the device) against a numpy restatement: kept
    keys, local column of every non-zero, owner segments, and slots that hold the right keys."""
    from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    dev = torch.device("cuda:0")
    data = make_rcv1_like(600, 5000, 12.0, seed=21, device=dev)
    for freq in (0, 1):
        cfg = AsyncSGDConfig(minibatch=300, tail_feature_freq=freq, countmin_n=1 << 24, table_capacity=1 << 14)
        eng = AsyncSGDEngine(cfg, None, dev)
        eng.load(data)
        for mb in range(2):
            m = eng._prepare(mb)
            lo, hi = int(data.row_ptr[mb * 300]), int(data.row_ptr[(mb + 1) * 300])
            keys = data.keys[lo:hi].cpu().numpy().view(np.uint64)
            kept = m.kept_keys().cpu().numpy().view(np.uint64)
            uniq, cnt = np.unique(keys, return_counts=True)
            if freq == 0:
                assert np.array_equal(kept, uniq)
            else:  # first minibatch: exactly the keys seen more than once (no collisions at this size)
                assert np.all(np.diff(kept.astype(np.int64)) > 0) and set(kept) <= set(uniq)
                if mb == 0:
                    assert np.array_equal(kept, uniq[cnt > freq])
            col = m.col[:hi - lo].cpu().numpy().view(np.uint32)
            pos = np.searchsorted(kept, keys)
            hit = (pos < kept.size) & (kept[np.minimum(pos, kept.size - 1)] == keys)
            assert np.array_equal(col[hit], pos[hit].astype(np.uint32)) and np.all(col[~hit] == 0xffffffff)
            assert m.owner_begin.tolist() == [0, kept.size]
            slots = m.col_slot[:kept.size].cpu().numpy()
            table = eng.b_keys.tensor.cpu().numpy().view(np.uint64)
            assert np.all(slots >= 0) and np.array_equal(table[slots], kept)
            assert np.all(m.col_owner[:kept.size].cpu().numpy() == 0)
        eng.close()
