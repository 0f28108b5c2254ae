"""Data-pipeline kernels (K10-K12, K15, K16) against plain torch references."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _keys(n, hi, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randint(0, hi, (n,), generator=g, device="cuda", dtype=torch.int64)


def test_unique_count_matches_torch():
    from parameter_server_b200.ops import localizer as L
    k = _keys(200000, 5000)
    u, c = L.unique_count(k)
    ur, cr = torch.unique(k, return_counts=True)
    assert torch.equal(u, ur) and torch.equal(c.long(), cr)


def test_unique_count_unsigned_order():
    from parameter_server_b200.ops import localizer as L
    # keys with the top bit set must sort AFTER small keys (uint64 order)
    k = torch.tensor([5, -1, 3, -(1 << 62), 5, 1 << 61], dtype=torch.int64, device="cuda")
    u, c = L.unique_count(k)
    as_u = [x & 0xFFFFFFFFFFFFFFFF for x in u.tolist()]
    assert as_u == sorted(as_u) and c.tolist() == [1, 2, 1, 1, 1]


def test_lookup_and_lower_bound():
    from parameter_server_b200.ops import localizer as L
    d = torch.unique(_keys(3000, 100000, 1))
    q = _keys(50000, 100000, 2)
    idx1 = L.lookup_sorted(d, q).long()
    pos = torch.searchsorted(d, q)
    hit = (pos < d.numel()) & (d[pos.clamp(max=d.numel() - 1)] == q)
    assert torch.equal(idx1, torch.where(hit, pos + 1, torch.zeros_like(pos)))
    assert torch.equal(L.lower_bound(d, q), pos)


def test_compact_and_transpose_roundtrip():
    from parameter_server_b200.ops import localizer as L
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    data = make_rcv1_like(4000, 900, 30.0, seed=2, device="cuda")
    dictionary, cnt = L.unique_count(data.keys)
    keep = dictionary[cnt > 3].contiguous()
    idx1 = L.lookup_sorted(keep, data.keys)
    rp, col, val = L.compact_csr(data.row_ptr, idx1, data.vals)
    m = idx1 != 0
    assert col.numel() == int(m.sum()) and int(rp[-1]) == col.numel()
    assert torch.equal(col.long(), idx1[m].long() - 1)
    assert torch.equal(val, data.vals[m])
    rows = torch.repeat_interleave(torch.arange(data.n_rows, device="cuda"),
                                   data.row_ptr[1:] - data.row_ptr[:-1])[m]
    assert torch.equal(torch.repeat_interleave(torch.arange(data.n_rows, device="cuda"),
                                               rp[1:] - rp[:-1]), rows)
    cp, ri, sv = L.csr_to_csc(rp, col, val, keep.numel(), data.y)
    dense = torch.zeros(data.n_rows, keep.numel(), dtype=torch.float64, device="cuda")
    dense[rows, col.long()] = (val * data.y[rows]).double()
    cols = torch.repeat_interleave(torch.arange(keep.numel(), device="cuda"), cp[1:] - cp[:-1])
    dense2 = torch.zeros_like(dense)
    dense2[ri.long(), cols] = sv.double()
    assert torch.equal(dense, dense2)
    # rows ascending inside every column (stable transpose)
    same_col = cols[1:] == cols[:-1]
    assert bool((ri.long()[1:][same_col] > ri.long()[:-1][same_col]).all())


def test_segments_cover_columns():
    from parameter_server_b200.ops import localizer as L
    nnz_per_col = torch.tensor([0, 5, 64, 65, 300, 1], device="cuda")
    cp = torch.zeros(7, dtype=torch.int64, device="cuda")
    cp[1:] = torch.cumsum(nnz_per_col, 0)
    col_seg, seg_col, seg_beg, seg_meta = L.build_segments(cp, 64)
    assert col_seg.tolist() == [0, 1, 2, 3, 5, 10, 11]
    assert seg_col.tolist() == [0, 1, 2, 3, 3, 4, 4, 4, 4, 4, 5]
    assert seg_beg.tolist()[3:5] == [int(cp[3]), int(cp[3]) + 64]
    m = seg_meta.view(-1, 4).tolist()
    assert [x[0] for x in m] == seg_col.tolist()
    assert [x[1] & 0xffff for x in m] == [0, 5, 64, 64, 1, 64, 64, 64, 64, 44, 1]
    assert [(x[1] >> 16) & 1 for x in m] == [1, 1, 1, 1, 0, 1, 0, 0, 0, 0, 1]
    assert [(x[1] >> 17) & 1 for x in m] == [1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 1]
    assert [x[2] for x in m] == seg_beg.tolist()


def test_key_signature_sensitive_to_content_and_order():
    from parameter_server_b200.ops import localizer as L
    k = torch.unique(_keys(10000, 1 << 40, 4))
    s0 = L.key_signature(k)
    assert s0 == L.key_signature(k.clone())
    k2 = k.clone(); k2[17] += 1
    assert L.key_signature(k2) != s0
    assert L.key_signature(torch.flip(k, [0])) != s0
