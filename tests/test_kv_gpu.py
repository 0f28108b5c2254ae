"""DeviceKVVector: push / pull through peer-memory kernels against plain torch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _check(world):
    from parameter_server_b200.parallel.comm import run_threads
    from parameter_server_b200.parallel.kv import DeviceKVVector
    dev = torch.device("cuda:0")
    dictionary = (torch.arange(50000, device=dev, dtype=torch.int64) * 7 + 3)

    def build(comm):
        return DeviceKVVector(dictionary, comm, dev, bulk_capacity=21000)

    kvs = run_threads(world, build)
    ref = torch.zeros(50000, dtype=torch.float64, device=dev)
    handles, vals = [], []
    for r, kv in enumerate(kvs):
        g = torch.Generator(device=dev).manual_seed(r)
        sel = torch.randperm(50000, generator=g, device=dev)[:20001].sort().values
        if r == 0:
            sel = torch.arange(1000, 21001, device=dev)  # a contiguous run exercises the 128-bit path
        h = kv.register(dictionary[sel])
        v = torch.randn(h.n, generator=g, device=dev)
        kv.push(h, v)
        kv.push(h, v)
        ref.index_add_(0, sel, 2 * v.double())
        handles.append((h, sel))
    torch.cuda.synchronize()
    assert sum(kv.shard_len[kv.rank] for kv in kvs) == 50000
    whole = torch.cat([kv.value for kv in kvs])
    assert torch.allclose(whole.double(), ref, atol=1e-5)
    for kv, (h, sel) in zip(kvs, handles):
        assert torch.allclose(kv.pull(h).double(), ref[sel], atol=1e-5)
        assert torch.equal(torch.cat([k.key for k in kvs]), dictionary)
    # bulk path, phase by phase (lock step): sends of every rank, then the owners' applies
    sels = []
    for r in range(world):
        g = torch.Generator(device=dev).manual_seed(50 + r)
        sels.append(torch.randperm(50000, generator=g, device=dev)[:20000].sort().values)

    def reg(comm):
        return kvs[comm.rank].register(dictionary[sels[comm.rank]], bulk=True)

    hs = run_threads(world, reg)
    vs = [torch.randn(h.n, device=dev) for h in hs]
    for kv, h, v in zip(kvs, hs, vs):
        kv._bulk_stage(h, v)     # every rank's values in its own staging buffer ...
    for kv, h in zip(kvs, hs):
        kv._push_apply(h)        # ... then every owner reads its segments from the "peers"
    for sel, v in zip(sels, vs):
        ref.index_add_(0, sel, v.double())
    torch.cuda.synchronize()
    assert torch.allclose(torch.cat([kv.value for kv in kvs]).double(), ref, atol=1e-5)
    for kv, h in zip(kvs, hs):
        kv._bulk_gather_send(h)
    for kv, h, sel in zip(kvs, hs, sels):
        assert torch.allclose(kv._bulk_pull_result(h).double(), ref[sel], atol=1e-5)
    for kv in kvs:
        kv.close()


def test_device_kv_single():
    _check(1)


def test_device_kv_three_emulated_ranks():
    _check(3)
