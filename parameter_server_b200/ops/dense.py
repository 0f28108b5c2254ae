"""Dense feature-block tile (dense_tile.cu): margins of n examples with d DENSE features against
K weight vectors on the tcgen05 tensor cores (TMA-fed, accumulator in TMEM), with the per-model
logistic loss reduced in the epilogue.  The device counterpart of the reference's generic SpMV on a
DENSE slot (src/data/text_parser.cc:200-250, src/util/sparse_matrix.h:72-124)."""
from __future__ import annotations

import ctypes as C

import torch

from .._lib import check, cuda_lib, current_stream, ptr


def _lib():
    lib = cuda_lib()
    if not getattr(lib, "_dense_sig", False):
        lib.psb_dense_margins_bf16.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
        lib.psb_dense_margins_bf16.restype = C.c_int
        lib._dense_sig = True
    return lib


def pad_features(x: torch.Tensor) -> torch.Tensor:
    """bf16 copy whose row stride is a multiple of 8 elements (16 bytes, what the tensor map needs)."""
    n, d = x.shape
    ld = (d + 7) // 8 * 8
    out = torch.zeros(n, ld, dtype=torch.bfloat16, device=x.device)
    out[:, :d] = x.to(torch.bfloat16)
    return out


def dense_margins(x: torch.Tensor, w: torch.Tensor, label: torch.Tensor, d: int | None = None,
                  with_loss: bool = True, with_margins: bool = True):
    """x: [n, ld] bf16, w: [K, ld'] bf16 (row strides multiples of 8), label: [n] float32 (+-1).
    Returns (Y [n, K] float32, loss [K] float64 or None): Y = x[:, :d] @ w[:, :d].T,
    loss[k] = sum_i log(1 + exp(-label_i Y[i, k]))."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.is_cuda and w.is_cuda
    assert x.stride(1) == 1 and w.stride(1) == 1
    n, k = x.shape[0], w.shape[0]
    d = d if d is not None else min(x.shape[1], w.shape[1])
    y = torch.empty(n, k, dtype=torch.float32, device=x.device) if with_margins else None
    loss = torch.zeros(k, dtype=torch.float64, device=x.device) if with_loss else None
    check(_lib().psb_dense_margins_bf16(ptr(x), x.stride(0), ptr(w), w.stride(0), ptr(y),
                                        ptr(label.float().contiguous()), ptr(loss), n, d, k, current_stream()))
    return y, loss
