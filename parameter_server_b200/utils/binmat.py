"""NumPy / SciPy counterparts of the reference's MATLAB helpers (src/data/matlab/*.m): raw
binary vectors, the row-major binary matrix format with its text `.info`, the PS text format
writer and the tail-feature filter.

    load_bin / save_bin        one typed vector per file
    mat2bin / bin2mat          {name}.index/.offset/.value/.label/.info  <->  (Y, X)
    saveas_pserver             "label; group id[:val] ...;" lines (ps sparse / sparse_binary)
    filter_fea                 drop columns seen <= pv times, keep the group offsets consistent
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp


def load_bin(filename: str, dtype="float64", offset: int = 0, length: int = -1) -> np.ndarray:
    dt = np.dtype(dtype)
    return np.fromfile(filename, dtype=dt, count=length, offset=offset * dt.itemsize)


def save_bin(filename: str, x, dtype="float64") -> None:
    np.ascontiguousarray(x, dtype=np.dtype(dtype)).tofile(filename)


def mat2bin(name: str, y, x) -> None:
    """Row-major binary matrix: uint32 column indices, uint64 row offsets, float64 values (omitted
    when every stored value is 1), float64 labels and a `key : value` info file."""
    x = sp.csr_matrix(x)
    x.sort_indices()
    save_bin(name + ".index", x.indices, "uint32")
    save_bin(name + ".offset", x.indptr, "uint64")
    binary = bool(np.all(x.data == 1))
    if not binary:
        save_bin(name + ".value", x.data, "float64")
    save_bin(name + ".label", np.asarray(y).ravel(), "float64")
    with open(name + ".info", "w") as f:
        f.write(f"row_major    : true\nsparse       : true\nbool_value   : {'true' if binary else 'false'}\n"
                f"row_begin    : 0\nrow_end      : {x.shape[0]}\ncol_begin    : 0\ncol_end      : {x.shape[1]}\n"
                f"entries      : {x.nnz}\nsizeof_index : 4\nsizeof_value : 8\n")


def _info(name: str) -> dict:
    out = {}
    for line in open(name + ".info"):
        if ":" in line:
            k, v = line.split(":", 1)
            out[k.strip()] = v.strip()
    return out


def bin2mat(name: str):
    """Inverse of mat2bin: returns (Y, X) with X a scipy CSR matrix."""
    info = _info(name)
    n_rows, n_cols = int(info["row_end"]) - int(info["row_begin"]), int(info["col_end"]) - int(info["col_begin"])
    indptr = load_bin(name + ".offset", "uint64").astype(np.int64)
    indices = load_bin(name + ".index", "uint32").astype(np.int64)
    if info.get("bool_value", "false") == "true":
        data = np.ones(indices.size, dtype=np.float64)
    else:
        data = load_bin(name + ".value", "float64")
    x = sp.csr_matrix((data, indices, indptr), shape=(n_rows, n_cols))
    y = load_bin(name + ".label", "float64")
    return y, x


def saveas_pserver(filename: str, y, x, group_id=None, binary: bool = False) -> None:
    """PS text: `label; gid key[:val] key[:val] ...; gid ...;` with keys local to the matrix."""
    x = sp.csr_matrix(x)
    x.sort_indices()
    gid = np.zeros(x.shape[1], dtype=np.int64) if group_id is None else np.asarray(group_id)
    assert np.all(np.diff(gid) >= 0), "group ids must be sorted by column"
    with open(filename, "w") as f:
        for i in range(x.shape[0]):
            lo, hi = x.indptr[i], x.indptr[i + 1]
            parts, prev = [f"{int(y[i])}"], None
            for j, v in zip(x.indices[lo:hi], x.data[lo:hi]):
                if gid[j] != prev:
                    prev = gid[j]
                    parts.append(f"; {prev}")
                parts.append(f" {j}" if binary else f" {j}:{v:g}")
            f.write("".join(parts) + ";\n")


def filter_fea(x, group_offset, pv: int = 4):
    """Keep the columns that occur in more than `pv` rows; returns (X', new group offsets)."""
    x = sp.csc_matrix(x)
    keep = np.diff(x.indptr) > pv
    off = np.asarray(group_offset, dtype=np.int64)
    kept_before = np.concatenate([[0], np.cumsum(keep)])
    return x[:, keep].tocsr(), kept_before[off]
