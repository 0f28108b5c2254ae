"""Device data-pipeline ops (K10-K12, K15, K16): thin torch wrappers over localizer.cu.

Keys are uint64 on the device.  torch has no usable uint64 arithmetic, so key tensors are
carried as int64 with the same bit pattern; every comparison happens inside the kernels
(unsigned), never in torch.

ref: src/util/localizer.h (countUniqIndex / remapIndex), src/util/sparse_matrix.h:185-241
"""
from __future__ import annotations

import ctypes

import torch

from .._lib import check, cuda_lib, current_stream, ptr
from ._sig import apply_cuda_sigs


def _lib():
    apply_cuda_sigs()
    return cuda_lib()


def _temp(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)


def _two_phase(fn, args_before, args_after, device):
    """Run a CUB-style entry: query temp size, allocate, run."""
    nbytes = ctypes.c_int64(0)
    check(fn(*args_before, None, ctypes.byref(nbytes), *args_after))
    tmp = _temp(nbytes.value, device)
    check(fn(*args_before, ptr(tmp), ctypes.byref(nbytes), *args_after))
    return tmp


def key_bits(max_key: int) -> int:
    return max(1, int(max_key).bit_length())


def sort_keys(keys: torch.Tensor, end_bit: int = 64) -> torch.Tensor:
    """Unsigned radix sort of a uint64 key tensor (stored as int64)."""
    assert keys.dtype == torch.int64 and keys.is_cuda
    out = torch.empty_like(keys)
    _two_phase(_lib().psb_sort_keys_u64, (ptr(keys), ptr(out), keys.numel(), 0, end_bit),
               (current_stream(),), keys.device)
    return out


def unique_count(keys: torch.Tensor, end_bit: int = 64):
    """K10: sorted unique keys and their multiplicities (uint32 stored as int32).

    ref: Localizer::countUniqIndex (src/util/localizer.h:69-108).  Counts are exact here;
    the reference's uint8 cast wraps at 256 — callers that feed a CountMin sketch saturate.
    """
    n = keys.numel()
    dev = keys.device
    if n == 0:
        return keys.new_empty(0), torch.empty(0, dtype=torch.int32, device=dev)
    srt = sort_keys(keys, end_bit)
    uniq = torch.empty_like(srt)
    cnt = torch.empty(n, dtype=torch.int32, device=dev)
    nruns = torch.zeros(1, dtype=torch.int64, device=dev)
    _two_phase(_lib().psb_rle_u64, (ptr(srt), n, ptr(uniq), ptr(cnt), ptr(nruns)),
               (current_stream(),), dev)
    m = int(nruns.item())
    return uniq[:m].clone(), cnt[:m].clone()


def lookup_sorted(dictionary: torch.Tensor, keys: torch.Tensor) -> torch.Tensor:
    """1 + position of each key in the sorted dictionary, 0 when absent (int32 bits of uint32)."""
    out = torch.empty(keys.numel(), dtype=torch.int32, device=keys.device)
    check(_lib().psb_lookup_sorted_u64(ptr(dictionary), dictionary.numel(), ptr(keys),
                                       keys.numel(), ptr(out), current_stream()))
    return out


def lower_bound(sorted_keys: torch.Tensor, probes: torch.Tensor) -> torch.Tensor:
    """K15: first position whose key is >= probe, for every probe (unsigned compare)."""
    out = torch.empty(probes.numel(), dtype=torch.int64, device=probes.device)
    check(_lib().psb_lower_bound_u64(ptr(sorted_keys), sorted_keys.numel(), ptr(probes),
                                     probes.numel(), ptr(out), current_stream()))
    return out


def compact_csr(row_ptr: torch.Tensor, idx1: torch.Tensor, val: torch.Tensor | None):
    """K11: drop entries with idx1==0, return (row_ptr', col uint32-as-int32, val')."""
    nnz = idx1.numel()
    n_rows = row_ptr.numel() - 1
    dev = idx1.device
    scan = torch.empty(max(nnz, 1), dtype=torch.int64, device=dev)
    col = torch.empty(nnz, dtype=torch.int32, device=dev)
    vout = torch.empty(nnz, dtype=torch.float32, device=dev) if val is not None else None
    lib = _lib()
    if nnz:
        _two_phase(lib.psb_compact_csr,
                   (ptr(row_ptr), n_rows, ptr(idx1), ptr(val), nnz, ptr(scan), None, ptr(col),
                    ptr(vout)), (current_stream(),), dev)
        nnz_new = int(scan[nnz - 1].item()) + int(idx1[nnz - 1].item() != 0)
    else:
        nnz_new = 0
    rp = torch.empty(n_rows + 1, dtype=torch.int64, device=dev)
    check(lib.psb_compact_row_ptr(ptr(row_ptr), n_rows, ptr(scan), nnz, nnz_new, ptr(rp),
                                  current_stream()))
    return rp, col[:nnz_new].clone(), (vout[:nnz_new].clone() if vout is not None else None)


def csr_to_csc(row_ptr: torch.Tensor, col: torch.Tensor, val: torch.Tensor | None, n_cols: int,
               y: torch.Tensor | None = None):
    """K12: stable CSR -> CSC.  With `y`, values are pre-multiplied by the row label."""
    nnz = col.numel()
    n_rows = row_ptr.numel() - 1
    dev = col.device
    col_ptr = torch.empty(n_cols + 1, dtype=torch.int64, device=dev)
    row_idx = torch.empty(nnz, dtype=torch.int32, device=dev)
    vout = torch.empty(nnz, dtype=torch.float32, device=dev) if val is not None else None
    work = torch.empty(3 * max(nnz, 1), dtype=torch.int32, device=dev)
    sorted_col = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
    _two_phase(_lib().psb_csr_to_csc,
               (ptr(row_ptr), ptr(col), ptr(val), ptr(y if val is not None else None), n_rows,
                n_cols, nnz, ptr(col_ptr), ptr(row_idx), ptr(vout), ptr(work), ptr(sorted_col)),
               (current_stream(),), dev)
    return col_ptr, row_idx, vout


def build_segments(col_ptr: torch.Tensor, seg_len: int):
    """Split every column into work segments of at most `seg_len` non-zeros."""
    n_cols = col_ptr.numel() - 1
    dev = col_ptr.device
    col_seg = torch.empty(n_cols + 1, dtype=torch.int64, device=dev)
    lib = _lib()
    _two_phase(lib.psb_segments_count, (ptr(col_ptr), n_cols, seg_len, ptr(col_seg)),
               (current_stream(),), dev)
    n_seg = int(col_seg[-1].item())
    seg_col = torch.empty(n_seg, dtype=torch.int32, device=dev)
    seg_beg = torch.empty(n_seg, dtype=torch.int64, device=dev)
    seg_meta = torch.empty(max(n_seg, 1) * 4, dtype=torch.int32, device=dev)
    check(lib.psb_segments_fill(ptr(col_ptr), ptr(col_seg), n_cols, seg_len, ptr(seg_col),
                                ptr(seg_beg), ptr(seg_meta), current_stream()))
    return col_seg, seg_col, seg_beg, seg_meta


def key_signature(keys: torch.Tensor) -> int:
    """K16: 64-bit signature of a key array (successor of the crc32c key-cache signature)."""
    out = torch.zeros(1, dtype=torch.int64, device=keys.device)
    check(_lib().psb_key_signature(ptr(keys), keys.numel(), ptr(out), current_stream()))
    return int(out.item()) & 0xFFFFFFFFFFFFFFFF
