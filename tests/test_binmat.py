"""utils/binmat.py: the NumPy counterparts of the reference's MATLAB data helpers."""
import numpy as np
import scipy.sparse as sp

from parameter_server_b200.runtime import host
from parameter_server_b200.utils import binmat


def test_bin_roundtrip_and_ps_text(tmp_path):
    rng = np.random.RandomState(0)
    x = sp.random(30, 12, density=0.3, random_state=rng, format="csr")
    y = np.where(rng.rand(30) > 0.5, 1.0, -1.0)
    binmat.mat2bin(str(tmp_path / "m"), y, x)
    y2, x2 = binmat.bin2mat(str(tmp_path / "m"))
    assert np.array_equal(y, y2) and (x != x2).nnz == 0
    xb = (x > 0).astype(np.float64)
    binmat.mat2bin(str(tmp_path / "b"), y, xb)
    assert not (tmp_path / "b.value").exists() and (binmat.bin2mat(str(tmp_path / "b"))[1] != xb).nnz == 0
    assert np.array_equal(binmat.load_bin(str(tmp_path / "m.label"), "float64", offset=3, length=4), y[3:7])
    # the PS text writer produces lines the native parser accepts (format SPARSE = 2)
    gid = np.array([1] * 5 + [2] * 7)
    binmat.saveas_pserver(str(tmp_path / "ps.txt"), y, x, gid)
    line = open(tmp_path / "ps.txt").readline().strip()
    exs, bad = host.parse_text(line, "sparse")
    assert bad == 0 and len(exs) == 1
    slots = {sid: (k, v) for sid, k, v in exs[0]}
    assert slots[0][1][0] == y[0]
    row = x[0].tocoo()
    got = np.concatenate([slots[g][0] for g in sorted(slots) if g != 0])
    assert np.array_equal(np.sort(got), np.sort(row.col.astype(np.uint64)))
    # tail filter keeps group boundaries consistent
    xf, off = binmat.filter_fea(x, [0, 5, 12], pv=8)
    cnt = np.diff(sp.csc_matrix(x).indptr)
    assert xf.shape[1] == int((cnt > 8).sum()) and off[0] == 0 and off[-1] == xf.shape[1]
    assert off[1] == int((cnt[:5] > 8).sum())
