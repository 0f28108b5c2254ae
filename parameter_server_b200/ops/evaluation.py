"""Device evaluation / sketch / codec ops (eval.cu): K9, K13, K14, K17-K19."""
from __future__ import annotations

import ctypes

import torch

from .._lib import check, cuda_lib, current_stream, ptr
from ._sig import apply_cuda_sigs


def _lib():
    apply_cuda_sigs()
    return cuda_lib()


def auc(label: torch.Tensor, predict: torch.Tensor) -> float:
    """Exact AUC of float32 predictions against +-1 labels (ref src/util/evaluation.h:22-48)."""
    n = label.numel()
    if n == 0:
        return float("nan")
    dev = label.device
    label = label.float().contiguous()
    predict = predict.float().contiguous()
    work_f = torch.empty(2 * n, dtype=torch.float32, device=dev)
    work_i = torch.empty(n, dtype=torch.int64, device=dev)
    out = torch.zeros(2, dtype=torch.float64, device=dev)
    nbytes = ctypes.c_int64(0)
    lib = _lib()
    check(lib.psb_auc_exact(ptr(label), ptr(predict), n, ptr(work_f), ptr(work_i), ptr(out), None,
                            ctypes.byref(nbytes), current_stream()))
    tmp = torch.empty(max(nbytes.value, 1), dtype=torch.uint8, device=dev)
    check(lib.psb_auc_exact(ptr(label), ptr(predict), n, ptr(work_f), ptr(work_i), ptr(out), ptr(tmp),
                            ctypes.byref(nbytes), current_stream()))
    area, pos = out.tolist()
    if pos == 0 or pos == n:
        return 1.0
    a = area / (pos * (n - pos))
    return 1 - a if a < 0.5 else a


def accuracy_logloss(label: torch.Tensor, predict: torch.Tensor, threshold: float = 0.0):
    n = label.numel()
    out = torch.zeros(2, dtype=torch.float64, device=label.device)
    check(_lib().psb_accuracy_logloss(ptr(label.float().contiguous()), ptr(predict.float().contiguous()), n,
                                      threshold, ptr(out), current_stream()))
    c, ll = out.tolist()
    acc = c / max(n, 1)
    return (acc if acc > 0.5 else 1 - acc), ll / max(n, 1)


class CountMin:
    """Device CountMin sketch, bit-compatible with the host sketch (same hash, probes, saturation)."""

    def __init__(self, n: int, k: int, device):
        self.n = max(64, (int(n) + 3) // 4 * 4)
        self.k = min(30, max(1, int(k)))
        self.table = torch.zeros(self.n, dtype=torch.uint8, device=device)

    def insert(self, keys: torch.Tensor, counts: torch.Tensor) -> None:
        check(_lib().psb_countmin_insert(ptr(keys), ptr(counts.to(torch.int32).contiguous()), keys.numel(),
                                         ptr(self.table), self.n, self.k, current_stream()))

    def query_gt(self, keys: torch.Tensor, freq: int) -> torch.Tensor:
        keep = torch.empty(keys.numel(), dtype=torch.uint8, device=keys.device)
        check(_lib().psb_countmin_query(ptr(keys), keys.numel(), ptr(self.table), self.n, self.k, freq,
                                        ptr(keep), current_stream()))
        return keep.bool()


def fixfloat_encode(x: torch.Tensor, nbytes: int, lo: float, hi: float, seed: int = 1) -> torch.Tensor:
    code = torch.empty(x.numel() * nbytes, dtype=torch.uint8, device=x.device)
    fn = _lib().psb_fixfloat_encode_f32 if x.dtype == torch.float32 else _lib().psb_fixfloat_encode_f64
    check(fn(ptr(x.contiguous()), x.numel(), nbytes, lo, hi, seed, ptr(code), current_stream()))
    return code


def fixfloat_decode(code: torch.Tensor, nbytes: int, lo: float, hi: float, dtype=torch.float32):
    n = code.numel() // nbytes
    out = torch.empty(n, dtype=dtype, device=code.device)
    fn = _lib().psb_fixfloat_decode_f32 if dtype == torch.float32 else _lib().psb_fixfloat_decode_f64
    check(fn(ptr(code), n, nbytes, lo, hi, ptr(out), current_stream()))
    return out


def fill_(t: torch.Tensor, kind: str = "ZERO", constant: float = 1.0, mean: float = 0.0, std: float = 1.0,
          seed: int = 1) -> torch.Tensor:
    assert t.dtype == torch.float64
    check(_lib().psb_fill_f64(ptr(t), t.numel(), {"ZERO": 1, "CONSTANT": 2, "GAUSSIAN": 3}[kind], constant,
                              mean, std, seed, current_stream()))
    return t


def bitmap_popcount(words: torch.Tensor) -> int:
    out = torch.zeros(1, dtype=torch.int64, device=words.device)
    check(_lib().psb_bitmap_popcount(ptr(words), words.numel(), ptr(out), current_stream()))
    return int(out.item())


def spmv_lookup(row_ptr, keys, val, dict_keys, dict_w) -> torch.Tensor:
    """Margins of CSR rows (64-bit keys) against a sorted (key, weight) model (K17)."""
    n_rows = row_ptr.numel() - 1
    out = torch.empty(max(n_rows, 1), dtype=torch.float32, device=keys.device)
    check(_lib().psb_spmv_lookup(ptr(row_ptr), ptr(keys), ptr(val), n_rows, ptr(dict_keys),
                                 ptr(dict_w.double().contiguous()), dict_keys.numel(), ptr(out),
                                 current_stream()))
    return out[:n_rows]
