"""Argument signatures of the C ABI (include/psb200.h), applied to the ctypes handles once.

Letters: p = pointer (void*), q = int64, Q = uint64, i = int32, d = double, f = float.
"""
from __future__ import annotations

import ctypes

from .._lib import cuda_lib

_T = {"p": ctypes.c_void_p, "q": ctypes.c_int64, "Q": ctypes.c_uint64, "i": ctypes.c_int32,
      "d": ctypes.c_double, "f": ctypes.c_float}

CUDA_SIGS = {
    # runtime.cu
    "psb_malloc": "pq", "psb_free": "p", "psb_memset_async": "piqp", "psb_memcpy_async": "ppqp",
    "psb_stream_sync": "p", "psb_host_alloc_pinned": "pq", "psb_host_free_pinned": "p",
    "psb_ipc_get_handle": "pp", "psb_ipc_open_handle": "pp", "psb_ipc_close_handle": "p",
    "psb_enable_peer_access": "i", "psb_bench_grid_barrier": "piiip", "psb_rank_barrier": "piiQp", "psb_rank_barrier_dev": "piipp",
    # localizer.cu
    "psb_sort_pairs_u64_u32": "ppppqiippp", "psb_sort_pairs_u32_u32": "ppppqiippp",
    "psb_sort_keys_u64": "ppqiippp", "psb_rle_u64": "pqpppppp", "psb_exclusive_sum_i64": "ppqppp",
    "psb_lookup_sorted_u64": "pqpqpp", "psb_lower_bound_u64": "pqpqpp",
    "psb_compact_csr": "pqppqppppppp", "psb_compact_row_ptr": "pqpqqpp",
    "psb_csr_to_csc": "ppppqqqpppppppp",
    "psb_segments_count": "pqipppp", "psb_segments_fill": "ppqipppp", "psb_key_signature": "pqpp",
    # eval.cu
    "psb_auc_exact": "ppqpppppp", "psb_accuracy_logloss": "ppqfpp",
    "psb_countmin_insert": "ppqpiip", "psb_countmin_query": "pqpiiipp",
    "psb_fixfloat_encode_f32": "pqiddQpp", "psb_fixfloat_decode_f32": "pqiddpp",
    "psb_fixfloat_encode_f64": "pqiddQpp", "psb_fixfloat_decode_f64": "pqiddpp",
    "psb_fill_f64": "pqidddQp", "psb_bitmap_popcount": "pqpp", "psb_spmv_lookup": "pppqppqpp",
    # kv.cu
    "psb_kv_route_dense": "pqpippp", "psb_kv_push_add_f32": "pppqpip", "psb_kv_pull_f32": "pppqpip",
    "psb_kv_bulk_send": "ppppiQpp", "psb_kv_bulk_apply_add": "pppqpQp",
    "psb_kv_bulk_gather_send": "pppppiQpp", "psb_kv_wait_flags": "piQp",
    "psb_kv_bulk_pull_apply": "ppppiipp", "psb_kv_bulk_gather_send32": "pppppiiQpp",
    # darlin.cu
    "psb_darlin_pass_begin": "pppp", "psb_darlin_grad_push": "pppiiip",
    "psb_darlin_server_update": "pppiiip", "psb_darlin_pull_update": "pppiiip",
    "psb_darlin_evaluate": "pppppiQpip", "psb_darlin_fused_pass": "pppppiiipppippp",
    "psb_darlin_fused_max_grid": "i", "psb_darlin_fused_pass_emulated": "piiiiiip", "psb_darlin_enqueue_pass": "ppppiiiippipip",
}
_applied = False


def apply_cuda_sigs() -> None:
    global _applied
    if _applied:
        return
    lib = cuda_lib()
    for name, sig in CUDA_SIGS.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = [_T[c] for c in sig]
        fn.restype = ctypes.c_int
    lib.psb_darlin_evaluate_scratch_bytes.argtypes = [ctypes.c_int32]
    lib.psb_darlin_evaluate_scratch_bytes.restype = ctypes.c_int64
    lib.psb_last_error.restype = ctypes.c_char_p
    _applied = True
