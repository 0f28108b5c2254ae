"""ctypes mirror of the Darlin structs of include/psb200.h plus call wrappers (darlin.cu)."""
from __future__ import annotations

import ctypes as C

from .._lib import check, cuda_lib, current_stream
from ._sig import apply_cuda_sigs

MAX_PEERS = 16
P = C.c_void_p


class DarlinCtrl(C.Structure):
    _fields_ = [("kkt_threshold", C.c_double), ("lambda_", C.c_double), ("eta", C.c_double),
                ("delta_max", C.c_double), ("seq_base", C.c_uint64),
                ("reset_kkt_filter", C.c_int32), ("n_steps", C.c_int32),
                ("revive_on_reset", C.c_int32), ("reserved0", C.c_int32),
                ("flags", C.c_int32), ("reserved", C.c_int32)]


class DarlinProgress(C.Structure):
    _fields_ = [("loss", C.c_double), ("penalty", C.c_double), ("violation", C.c_double),
                ("nnz_w", C.c_uint64), ("nnz_active", C.c_uint64), ("seq", C.c_uint64)]


class DarlinWorker(C.Structure):
    _fields_ = [("col_ptr", P), ("row_idx", P), ("sval", P), ("y", P), ("z", P), ("w", P),
                ("delta", P), ("active", P), ("col_owner", P), ("col_off", P), ("seg_col", P),
                ("seg_beg", P), ("col_seg", P), ("blk_col", P), ("blk_srv_mask", P), ("part", P),
                ("col_cnt", P), ("done_ctr", P),
                ("n_rows", C.c_int64), ("n_cols", C.c_int64), ("seg_len", C.c_int32),
                ("rank", C.c_int32), ("n_workers", C.c_int32), ("n_servers", C.c_int32),
                ("land_g", P * MAX_PEERS), ("land_u", P * MAX_PEERS),
                ("land_cap", C.c_int64 * MAX_PEERS), ("push_flag", P * MAX_PEERS),
                ("srv_w", P * MAX_PEERS), ("srv_dw", P * MAX_PEERS), ("ver_flag", P),
                ("seg_meta", P), ("land2", P * MAX_PEERS), ("col_has", P), ("stash", P),
                ("done_local", P), ("done_peer", P * MAX_PEERS), ("push_fp32", C.c_int32),
                ("reserved1", C.c_int32), ("acc", P)]


class DarlinServer(C.Structure):
    _fields_ = [("w", P), ("dw", P), ("delta", P), ("active", P), ("land_g", P), ("land_u", P),
                ("land_cap", C.c_int64), ("push_flag", P), ("blk_rng", P), ("violation", P),
                ("done_ctr", P), ("n_keys", C.c_int64), ("rank", C.c_int32),
                ("n_workers", C.c_int32), ("n_servers", C.c_int32),
                ("ver_flag", P * MAX_PEERS)]


class DarlinRankArgs(C.Structure):
    """psb_darlin_rank_args: one rank of the single-GPU emulation of the multi-rank pass."""
    _fields_ = [("wk", DarlinWorker), ("sv", DarlinServer), ("ctrl", P), ("order", P),
                ("blk_tab", P), ("bar", P), ("out_local", P), ("out_peer", P), ("scratch", P)]


def _lib():
    apply_cuda_sigs()
    return cuda_lib()


def _ref(s):
    return C.byref(s) if s is not None else None


def pass_begin(wk, sv, ctrl_ptr: int):
    check(_lib().psb_darlin_pass_begin(_ref(wk), _ref(sv), ctrl_ptr, current_stream()))


def grad_push(wk, ctrl_ptr: int, order_ptr: int, step: int, bps: int, grid: int):
    check(_lib().psb_darlin_grad_push(_ref(wk), ctrl_ptr, order_ptr, step, bps, grid,
                                      current_stream()))


def server_update(sv, ctrl_ptr: int, order_ptr: int, step: int, bps: int, grid: int):
    check(_lib().psb_darlin_server_update(_ref(sv), ctrl_ptr, order_ptr, step, bps, grid,
                                          current_stream()))


def pull_update(wk, ctrl_ptr: int, order_ptr: int, step: int, bps: int, grid: int):
    check(_lib().psb_darlin_pull_update(_ref(wk), ctrl_ptr, order_ptr, step, bps, grid,
                                        current_stream()))


def evaluate(wk, sv, ctrl_ptr: int, out_local: int, out_peer, n_peers: int, seq: int,
             scratch_ptr: int, grid: int):
    check(_lib().psb_darlin_evaluate(_ref(wk), _ref(sv), ctrl_ptr, out_local, out_peer, n_peers,
                                     seq, scratch_ptr, grid, current_stream()))


def evaluate_scratch_bytes(grid: int) -> int:
    return int(_lib().psb_darlin_evaluate_scratch_bytes(grid))


def enqueue_pass(wk, sv, ctrl_ptr: int, order_ptr: int, n_steps: int, bps: int, grid_w: int,
                 grid_s: int, out_local: int, out_peer, n_peers: int, scratch_ptr: int,
                 grid_eval: int):
    check(_lib().psb_darlin_enqueue_pass(_ref(wk), _ref(sv), ctrl_ptr, order_ptr, n_steps, bps,
                                         grid_w, grid_s, out_local, out_peer, n_peers,
                                         scratch_ptr, grid_eval, current_stream()))


def fused_max_grid(binary: bool) -> int:
    return int(_lib().psb_darlin_fused_max_grid(int(binary)))


def fused_pass(wk, sv, ctrl_ptr: int, order_ptr: int, blk_tab_ptr: int, n_steps: int, bps: int,
               grid: int, bar_ptr: int, out_local: int, out_peer, n_peers: int, scratch_ptr: int,
               prof_ptr=None):
    check(_lib().psb_darlin_fused_pass(_ref(wk), _ref(sv), ctrl_ptr, order_ptr, blk_tab_ptr,
                                       n_steps, bps, grid, bar_ptr, out_local, out_peer, n_peers,
                                       scratch_ptr, prof_ptr, current_stream()))


def fused_pass_emulated(ranks_dev_ptr: int, n_ranks: int, binary: bool, n_steps: int, bps: int,
                        grid_per_rank: int, n_peers: int):
    check(_lib().psb_darlin_fused_pass_emulated(ranks_dev_ptr, n_ranks, int(binary), n_steps, bps,
                                                grid_per_rank, n_peers, current_stream()))
