"""Asynchronous minibatch SGD (FTRL / AdaGrad / SGD) on B200: every rank is a worker streaming
minibatches and a server shard holding an open-addressing table {key, w, z, n} in HBM.

ref: src/app/linear_method/async_sgd.h (scheduler / server / worker), src/learner/sgd.h
(MinibatchReader), src/parameter/kv_store.h.  Per minibatch the worker runs ONE cooperative
kernel (pull weights from the owners' tables in peer memory -> margins, loss, gradient -> push
(slot, g) into the owners' landing rings, optionally int8 block-scaled), the server applies the
landed pushes in worker order and acknowledges.  `max_delay` = tau is honoured: a worker owns
tau+1 landing regions per server and may run at most tau pushes ahead of the acknowledgements.

Two ways to feed it:
  load(data) + train()      the whole data set resident in HBM; the (owner, slot) resolution of
                            a minibatch is cached for the following passes up to `cache_bytes`
                            (the device successor of the KEY_CACHING filter);
  train_stream(source)      minibatches arrive from the host (the StreamReader of the reference,
                            src/learner/sgd.h:104-157): the native runner (csrc/cuda/sgd_stream.cu)
                            copies them from pinned memory into a ring of device buffers and
                            localises them on a copy stream while the step kernels of earlier
                            minibatches run; the interpreter only hands over windows of
                            minibatch descriptors, and nothing is kept.
The tail-feature filter is the reference's: one CountMin sketch per worker that lives across
minibatches (insert the counts of a minibatch, keep the keys whose estimate exceeds the bound).
"""
from __future__ import annotations

import ctypes as C
import math
import os
import sys
import time
from dataclasses import dataclass

import numpy as np
import torch

from .._lib import check, cuda_lib, current_stream, ptr
from ..ops._sig import apply_cuda_sigs
from ..parallel.comm import Comm, SingleComm
from ..utils.range import Range
from ..utils.synthetic import SparseCSR

MAX_PEERS = 16
P = C.c_void_p
LOSS = {"SQUARE": 1, "LOGIT": 2, "SQUARE_HINGE": 4}
ALGO = {"FTRL": 0, "ADAGRAD": 1, "SGD": 2}


class SgdWorkerStruct(C.Structure):
    _fields_ = [("row_ptr", P), ("col", P), ("val", P), ("y", P), ("n_rows", C.c_int64),
                ("n_cols", C.c_int64), ("col_owner", P), ("col_slot", P), ("owner_begin", P),
                ("w_local", P), ("g_local", P), ("xw", P), ("stats", P),
                ("tab_w", P * MAX_PEERS), ("land_slot", P * MAX_PEERS), ("land_val", P * MAX_PEERS),
                ("land_q", P * MAX_PEERS), ("land_scale", P * MAX_PEERS), ("land_cnt", P * MAX_PEERS),
                ("land_flag", P * MAX_PEERS), ("ack", P), ("land_cap", C.c_int64),
                ("rank", C.c_int32), ("n_workers", C.c_int32), ("n_servers", C.c_int32),
                ("depth", C.c_int32), ("loss", C.c_int32), ("quant", C.c_int32), ("n_cols_dev", P)]


class SgdServerStruct(C.Structure):
    _fields_ = [("w", P), ("z", P), ("n", P), ("cap", C.c_int64), ("land_slot", P), ("land_val", P),
                ("land_q", P), ("land_scale", P), ("land_cnt", P), ("land_flag", P),
                ("ack_out", P * MAX_PEERS), ("land_cap", C.c_int64), ("stats", P),
                ("rank", C.c_int32), ("n_workers", C.c_int32), ("depth", C.c_int32),
                ("algo", C.c_int32), ("quant", C.c_int32), ("alpha", C.c_float), ("beta", C.c_float),
                ("l1", C.c_float), ("l2", C.c_float), ("sgd_t", C.c_float), ("lr_constant", C.c_int32)]


MAX_RING = 12
MAX_LANES = 8


class HostMinibatchStruct(C.Structure):
    _fields_ = [("row_ptr", P), ("keys", P), ("vals", P), ("y", P), ("n_rows", C.c_int64), ("nnz", C.c_int64)]


class RingSetStruct(C.Structure):
    _fields_ = [("row_ptr", P), ("keys", P), ("vals", P), ("y", P), ("cols", P), ("col", P), ("col_owner", P),
                ("col_slot", P), ("counts", P), ("owner_begin", P)]


class ScratchStruct(C.Structure):
    _fields_ = [("sorted", P), ("uniq", P), ("cnt", P), ("flag", P), ("temp", P)]


class RunnerConfigStruct(C.Structure):
    _fields_ = [("worker", SgdWorkerStruct), ("server", SgdServerStruct), ("set", RingSetStruct * MAX_RING),
                ("n_sets", C.c_int32), ("fused_server", C.c_int32), ("max_rows", C.c_int64), ("max_nnz", C.c_int64),
                ("scratch", ScratchStruct * MAX_LANES), ("temp_bytes", C.c_int64),
                ("sketch", P), ("cm_n", C.c_int32), ("cm_k", C.c_int32), ("freq", C.c_int32), ("reserved", C.c_int32),
                ("bounds_dev", P), ("table_keys", P * MAX_PEERS), ("table_cap", C.c_int64 * MAX_PEERS),
                ("overflow_dev", P), ("overflow_host", P), ("worker_bar", P), ("server_bar", P),
                ("grid_worker", C.c_int32), ("grid_server", C.c_int32), ("copy_stream", P * MAX_LANES),
                ("n_lanes", C.c_int32), ("reserved2", C.c_int32)]


@dataclass
class AsyncSGDConfig:
    """LM::Config fields that drive async SGD (linear.proto: SGDConfig, penalty, learning_rate)."""
    algo: str = "FTRL"               # FTRL | ADAGRAD | SGD
    loss: str = "LOGIT"
    minibatch: int = 1000
    num_data_pass: int = 1
    max_delay: int = 0               # tau
    alpha: float = 1.0
    beta: float = 1.0
    lambda_l1: float = 0.0
    lambda_l2: float = 0.0
    lr_type: str = "DECAY"           # learning_rate.type: DECAY alpha/(x+beta) | CONSTANT alpha
    tail_feature_freq: int = 0
    countmin_n: int = 100_000_000    # cells / probes of the worker's tail-filter sketch (linear.proto:60-61)
    countmin_k: int = 2
    fixing_float_by_nbytes: int = 0  # != 0 -> block-scaled 8-bit push (K9), see push_format
    push_format: str = "int8"        # int8 | fp8 (e4m3), one fp32 scale per 128 values
    cache_bytes: int = 8 << 30       # resident mode: budget for cached localised minibatches
    prefetch: int = 3                # streaming mode: localised minibatches kept ahead of the step (ring = +1)
    stream_window: int = 64          # streaming mode: minibatches handed to the native runner per call
    copy_lanes: int = 4              # streaming mode: minibatches uploaded / localised concurrently (1..8)
    table_capacity: int = 1 << 22    # slots per shard (power of two)
    report_every: int = 10           # minibatches between progress lines
    fused_server: bool = True        # one cooperative server launch per step instead of 1 + n_workers

    @staticmethod
    def from_proto_text(conf_text: str) -> "AsyncSGDConfig":
        from ..runtime.host import textproto_get as g
        c = AsyncSGDConfig()
        base = "linear_method."
        algo = g(conf_text, base + "async_sgd.algo") or "FTRL"
        ada = (g(conf_text, base + "async_sgd.ada_grad") or "true") == "true"
        c.algo = "FTRL" if algo == "FTRL" else ("ADAGRAD" if ada else "SGD")
        c.loss = g(conf_text, base + "loss.type") or "LOGIT"
        for name, typ in (("minibatch", int), ("num_data_pass", int), ("max_delay", int),
                          ("tail_feature_freq", int), ("fixing_float_by_nbytes", int),
                          ("countmin_n", int), ("countmin_k", int)):
            v = g(conf_text, base + "async_sgd." + name)
            if v:
                setattr(c, name, typ(float(v)))
        c.alpha = float(g(conf_text, base + "learning_rate.alpha") or 1)
        c.beta = float(g(conf_text, base + "learning_rate.beta") or 1)
        c.lr_type = g(conf_text, base + "learning_rate.type") or "DECAY"
        ptype = g(conf_text, base + "penalty.type") or "L1"
        l0 = float(g(conf_text, base + "penalty.lambda[0]") or 0)
        l1 = float(g(conf_text, base + "penalty.lambda[1]") or 0)
        c.lambda_l1, c.lambda_l2 = (l0, l1) if ptype == "L1" else (0.0, l0)
        return c


@dataclass
class _Minibatch:
    row_ptr: torch.Tensor
    col: torch.Tensor
    val: torch.Tensor | None
    y: torch.Tensor
    keys: torch.Tensor
    col_owner: torch.Tensor
    col_slot: torch.Tensor
    owner_begin: torch.Tensor
    n_rows: int
    n_cols: int                       # capacity of the column arrays (the non-zeros of the minibatch)
    n_cols_dev: torch.Tensor | None = None   # [1] int64 on the device: the actual number of columns
    xw: torch.Tensor | None = None           # streaming: the buffer that holds this step's margins afterwards

    def kept_keys(self) -> torch.Tensor:
        """The minibatch's columns (sorted unique keys that passed the tail filter); synchronises."""
        return self.keys[:int(self.n_cols_dev.item())] if self.n_cols_dev is not None else self.keys[:self.n_cols]


def _lib():
    apply_cuda_sigs()
    lib = cuda_lib()
    if not getattr(lib, "_sgd_sigs", False):
        lib.psb_kv_table_clear.argtypes = [P, C.c_int64, P]
        lib.psb_kv_lookup_insert.argtypes = [P, C.c_int64, P, P, P, C.c_int32, P, P, P]
        lib.psb_kv_export_nonzero.argtypes = [P, P, C.c_int64, P, P, P, P]
        lib.psb_sgd_worker_step.argtypes = [P, C.c_uint64, C.c_int32, P, P]
        lib.psb_sgd_server_update.argtypes = [P, C.c_uint64, C.c_int32, P]
        lib.psb_sgd_server_update_fused.argtypes = [P, C.c_uint64, C.c_int32, P, P]
        lib.psb_sgd_max_grid.restype = C.c_int
        lib.psb_sgd_set_fault_sink.argtypes = [P]
        lib.psb_sgd_localize_temp_bytes.argtypes = [C.c_int64]
        lib.psb_sgd_localize_temp_bytes.restype = C.c_int64
        lib.psb_sgd_runner_create.argtypes = [P]
        lib.psb_sgd_runner_create.restype = P
        lib.psb_sgd_runner_destroy.argtypes = [P]
        lib.psb_sgd_runner_destroy.restype = None
        lib.psb_sgd_runner_timing.argtypes = [P, P]
        lib.psb_sgd_runner_submit.argtypes = [P, P, C.c_int32, C.c_uint64, P]
        lib.psb_sgd_localize.argtypes = [P, C.c_int64, P, P, P, P, P, P, P, P, C.c_int32, C.c_int32, C.c_int32,
                                         P, C.c_int32, C.c_int64, P, P, P, P, P, P, P, C.c_int64, P]
        lib._sgd_sigs = True
    return lib


class AsyncSGDEngine:
    def __init__(self, config: AsyncSGDConfig, comm: Comm | None = None, device=None, out=sys.stderr):
        self.cfg = config
        self.comm = comm if comm is not None else SingleComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        self.device = torch.device(device if device is not None else
                                   f"cuda:{torch.cuda.current_device()}")
        self.out = out
        self.lib = _lib()
        self.seq = 0
        self.history = []   # (examples, loss, accuracy, auc)
        self._cache = {}
        self._cache_used = 0
        self._scratches = {}
        self._wk = self._w_local = self._g_local = self._xw = None
        self._step_launches = 0   # written by the step loop
        self._loc_launches = 0    # written by whichever thread localises
        self.h2d_bytes = 0
        self.submit_s = 0.0          # streaming: wall time inside the native runner (enqueueing)
        self.source_s = 0.0          # streaming: wall time spent pulling minibatches from the source
        self._runner = None
        self.data = None
        self.n_rows = 0
        # PSB_SGD_PROFILE=1: wall time per localisation stage (synchronising after each: slower)
        self._stage_prof = {} if os.environ.get("PSB_SGD_PROFILE") else None

    @property
    def launches(self) -> int:
        """Kernels of this framework launched so far (step kernels + localisation kernels)."""
        return self._step_launches + self._loc_launches

    # --------------------------------------------------------------------------------------------
    def load(self, data: SparseCSR, max_unique_per_minibatch: int | None = None) -> None:
        """Resident mode: the worker's rows live in HBM for all passes."""
        dev = self.device
        self.data = data.to(dev)
        self.n_rows = data.n_rows
        max_row = int((self.data.row_ptr[1:] - self.data.row_ptr[:-1]).max().item()) if self.n_rows else 1
        cap = max_unique_per_minibatch or min(self.cfg.minibatch * max(max_row, 1), self.data.nnz + 1)
        self.open(cap)
        self.n_minibatches = (self.n_rows + self.cfg.minibatch - 1) // self.cfg.minibatch
        # every rank runs the same number of steps (ranks with fewer rows send empty pushes)
        self.steps_per_pass = int(max(self.comm.gather_obj(self.n_minibatches)))

    def open(self, max_unique_per_minibatch: int) -> None:
        """Allocate the shard (hash table + optimizer state), the landing rings and the flags.
        `max_unique_per_minibatch`: upper bound of the distinct keys one minibatch sends to one
        owner (the size of a ring region)."""
        cfg, dev, W = self.cfg, self.device, self.world
        assert cfg.table_capacity & (cfg.table_capacity - 1) == 0
        self.depth = max(1, cfg.max_delay + 1)
        # the arena layout must be identical on every rank: take the largest request
        land_cap = int(max(self.comm.gather_obj(int(max_unique_per_minibatch))))
        land_cap = (land_cap + 127) // 128 * 128
        self.land_cap = land_cap
        nblk = land_cap // 128 + 1
        cap = cfg.table_capacity
        D = self.depth
        nbytes = cap * (8 + 4 * 3) + W * D * land_cap * (4 + 4 + 1) + W * D * nblk * 4 + \
            (W * D + 3 * W + 16) * 8 + 64 * 256
        self.arena = self.comm.make_arena(nbytes, dev)
        A = self.arena
        self.b_keys = A.alloc("tab_keys", cap, torch.int64)
        self.b_w = A.alloc("tab_w", cap, torch.float32)
        self.b_z = A.alloc("tab_z", cap, torch.float32)
        self.b_n = A.alloc("tab_n", cap, torch.float32)
        self.b_lslot = A.alloc("land_slot", W * D * land_cap, torch.int32)
        self.b_lval = A.alloc("land_val", W * D * land_cap, torch.float32)
        self.b_lq = A.alloc("land_q", W * D * land_cap, torch.int8)
        self.b_lscale = A.alloc("land_scale", W * D * nblk, torch.float32)
        self.b_lcnt = A.alloc("land_cnt", W * D, torch.int64)
        self.b_lflag = A.alloc("land_flag", W, torch.int64)
        self.b_ack = A.alloc("ack", W, torch.int64)
        self.b_bar = A.alloc("bar_flag", W, torch.int64)
        check(self.lib.psb_kv_table_clear(self.b_keys.ptr, cap, current_stream()))
        torch.cuda.synchronize(dev)
        self.comm.barrier()
        f64 = dict(dtype=torch.float64, device=dev)
        self.wstats = torch.zeros(4, **f64)
        self.sstats = torch.zeros(4, **f64)
        self.bar = torch.zeros(4, dtype=torch.int32, device=dev)
        self.srv_bar = torch.zeros(4, dtype=torch.int32, device=dev)
        self.overflow = torch.zeros(1, dtype=torch.int32, device=dev)
        self.grid = int(self.lib.psb_sgd_max_grid())
        # device waits that time out report here instead of hanging the GPU (sgd.cu spin_or_fault)
        self.fault = torch.zeros(4, dtype=torch.int64).pin_memory()
        check(self.lib.psb_sgd_set_fault_sink(self.fault.data_ptr()))
        sm = torch.cuda.get_device_properties(dev).multi_processor_count
        self.grid_server = sm * 2
        whole = Range.all()
        self.server_ranges = [whole.even_divide(W, s) for s in range(W)]
        self.bounds = torch.tensor([_as_i64(r.begin) for r in self.server_ranges[1:]],
                                   dtype=torch.int64, device=dev)
        self.tab_keys_ptrs = (P * W)(*[A.peer_ptr(s, self.b_keys) for s in range(W)])
        self.tab_caps = (C.c_int64 * W)(*[cap] * W)
        self.quant = 0 if not cfg.fixing_float_by_nbytes else {"int8": 1, "fp8": 2}[cfg.push_format]
        sv = SgdServerStruct()
        sv.w, sv.z, sv.n, sv.cap = self.b_w.ptr, self.b_z.ptr, self.b_n.ptr, cap
        sv.land_slot, sv.land_val, sv.land_q = self.b_lslot.ptr, self.b_lval.ptr, self.b_lq.ptr
        sv.land_scale, sv.land_cnt, sv.land_flag = self.b_lscale.ptr, self.b_lcnt.ptr, self.b_lflag.ptr
        for w in range(W):
            sv.ack_out[w] = A.peer_ptr(w, self.b_ack, self.rank)
        sv.land_cap, sv.stats = land_cap, self.sstats.data_ptr()
        sv.rank, sv.n_workers, sv.depth = self.rank, W, D
        sv.algo, sv.quant = ALGO[cfg.algo], self.quant
        sv.alpha, sv.beta, sv.l1, sv.l2, sv.sgd_t = cfg.alpha, cfg.beta, cfg.lambda_l1, cfg.lambda_l2, 1.0
        sv.lr_constant = int(cfg.lr_type.upper() == "CONSTANT")
        self.sv = sv
        # the worker's tail-feature sketch lives across minibatches and passes (ref sgd.h:137-147)
        self.tail_sketch = None
        if cfg.tail_feature_freq > 0:
            from ..ops.evaluation import CountMin
            self.tail_sketch = CountMin(min(cfg.countmin_n, 1 << 30), cfg.countmin_k, dev)

    # --------------------------------------------------------------------------------------------
    def _prepare(self, mb: int) -> _Minibatch:
        """Resident mode: localise minibatch `mb` of the loaded data (cached across passes while
        the budget lasts; never with a tail filter, whose answer changes as counts accumulate)."""
        if mb in self._cache:
            return self._cache[mb]
        cfg, d = self.cfg, self.data
        a = min(mb * cfg.minibatch, self.n_rows)
        b = min(a + cfg.minibatch, self.n_rows)
        lo, hi = int(d.row_ptr[a].item()), int(d.row_ptr[b].item())
        m = self._localize((d.row_ptr[a:b + 1] - lo).contiguous(), d.keys[lo:hi].contiguous(),
                           d.vals[lo:hi].contiguous() if d.vals is not None else None, d.y[a:b].contiguous())
        size = sum(t.numel() * t.element_size() for t in (m.row_ptr, m.col, m.y, m.keys, m.col_owner, m.col_slot)
                   ) + (m.val.numel() * 4 if m.val is not None else 0)
        if self.tail_sketch is None and self._cache_used + size <= cfg.cache_bytes:
            self._cache[mb] = m
            self._cache_used += size
        return m

    def _localize(self, row_ptr: torch.Tensor, keys_nnz: torch.Tensor, vals, y: torch.Tensor) -> _Minibatch:
        """K10 + K13 + K11 + K15 + K8a as ONE enqueue on the current stream (psb_sgd_localize):
        unique keys and counts, tail filter, local column of every non-zero, owner segments, slots
        in the owners' tables.  No size that depends on the data comes back to the host: the
        number of columns stays in `n_cols_dev`, and the step kernel reads it there.  The arrays of
        the returned minibatch have the capacity of its non-zeros."""
        cfg, dev, W = self.cfg, self.device, self.world
        t0 = time.perf_counter() if self._stage_prof is not None else 0.0
        nnz, n_rows = keys_nnz.numel(), row_ptr.numel() - 1
        cap = max(nnz, 1)
        i64 = dict(dtype=torch.int64, device=dev)
        cols = torch.empty(cap, **i64)
        col_slot = torch.empty(cap, **i64)
        col = torch.empty(cap, dtype=torch.int32, device=dev)
        col_owner = torch.empty(cap, dtype=torch.int32, device=dev)
        counts = torch.empty(2, **i64)
        owner_begin = torch.empty(W + 1, **i64)
        sc = self._scratch(cap)
        sk = self.tail_sketch
        check(self.lib.psb_sgd_localize(
            ptr(keys_nnz), nnz, ptr(sc["sorted"]), ptr(sc["uniq"]), ptr(sc["cnt"]), ptr(sc["flag"]),
            ptr(cols), ptr(col), ptr(counts),
            sk.table.data_ptr() if sk is not None else None, sk.n if sk is not None else 0,
            sk.k if sk is not None else 0, cfg.tail_feature_freq,
            ptr(self.bounds), W, self.land_cap, ptr(owner_begin), self.tab_keys_ptrs, self.tab_caps,
            ptr(col_owner), ptr(col_slot), ptr(self.overflow), ptr(sc["temp"]), sc["temp"].numel(),
            current_stream()))
        self._loc_launches += 5 if sk is not None else 4   # own kernels; the sort / run-length / select are CUB's
        self._new_minibatches = getattr(self, "_new_minibatches", 0) + 1
        if self._new_minibatches % 64 == 0:
            self.check_table_load()
        if self._stage_prof is not None:
            torch.cuda.current_stream(dev).synchronize()
            self._stage_prof["localize"] = self._stage_prof.get("localize", 0.0) + time.perf_counter() - t0
        return _Minibatch(row_ptr, col, vals, y, cols, col_owner, col_slot, owner_begin, n_rows, nnz, counts[1:])

    def _scratch(self, nnz: int) -> dict:
        """Work arrays of the localisation, one set per stream that localises (the prefetch stream in
        streaming mode, the step stream in resident mode): calls on one stream are ordered, so the set
        is reused from minibatch to minibatch."""
        key = torch.cuda.current_stream(self.device).cuda_stream
        sc = self._scratches.get(key)
        if sc is None or sc["cap"] < nnz:
            cap = max(nnz, 2 * sc["cap"] if sc else 0)
            dev = self.device
            sc = dict(cap=cap,
                      sorted=torch.empty(cap, dtype=torch.int64, device=dev),
                      uniq=torch.empty(cap, dtype=torch.int64, device=dev),
                      cnt=torch.empty(cap, dtype=torch.int32, device=dev),
                      flag=torch.empty(cap, dtype=torch.uint8, device=dev),
                      temp=torch.empty(int(self.lib.psb_sgd_localize_temp_bytes(cap)), dtype=torch.uint8, device=dev))
            self._scratches[key] = sc
        return sc

    def _worker_template(self, max_cols: int, max_rows: int):
        """The part of the step's launch argument that does not depend on the minibatch: the peer
        views (filled in once) and the buffers of the pulled weights, the gradient and the margins,
        grown to at least (max_cols, max_rows) and reused from step to step - all steps run on one
        stream."""
        wk = self._wk
        if wk is None:
            cfg, A, W, D, r = self.cfg, self.arena, self.world, self.depth, self.rank
            nblk = self.land_cap // 128 + 1
            wk = SgdWorkerStruct()
            wk.stats = self.wstats.data_ptr()
            for s in range(W):
                wk.tab_w[s] = A.peer_ptr(s, self.b_w)
                wk.land_slot[s] = A.peer_ptr(s, self.b_lslot, r * D * self.land_cap)
                wk.land_val[s] = A.peer_ptr(s, self.b_lval, r * D * self.land_cap)
                wk.land_q[s] = A.peer_ptr(s, self.b_lq, r * D * self.land_cap)
                wk.land_scale[s] = A.peer_ptr(s, self.b_lscale, r * D * nblk)
                wk.land_cnt[s] = A.peer_ptr(s, self.b_lcnt, r * D)
                wk.land_flag[s] = A.peer_ptr(s, self.b_lflag, r)
            wk.ack = self.b_ack.ptr
            wk.land_cap = self.land_cap
            wk.rank, wk.n_workers, wk.n_servers, wk.depth = r, W, W, D
            wk.loss, wk.quant = LOSS[cfg.loss], self.quant
            self._wk = wk
        have_c = self._w_local.numel() if self._w_local is not None else 0
        have_r = self._xw.numel() if self._xw is not None else 0
        if have_c < max_cols or have_r < max_rows:
            dev = self.device
            nc, nr = max(max_cols, 2 * have_c, 1), max(max_rows, have_r, 1)
            self._w_local = torch.empty(nc, dtype=torch.float32, device=dev)
            self._g_local = torch.empty(nc, dtype=torch.float32, device=dev)
            self._xw = torch.empty(nr, dtype=torch.float32, device=dev)
            wk.w_local, wk.g_local, wk.xw = ptr(self._w_local), ptr(self._g_local), ptr(self._xw)
        return wk

    def _worker_struct(self, m: _Minibatch):
        """The launch argument of the step on minibatch `m`."""
        wk = self._worker_template(m.n_cols, m.n_rows)
        wk.row_ptr, wk.col, wk.val, wk.y = ptr(m.row_ptr), ptr(m.col), ptr(m.val), ptr(m.y)
        wk.n_rows, wk.n_cols = m.n_rows, m.n_cols
        wk.col_owner, wk.col_slot, wk.owner_begin = ptr(m.col_owner), ptr(m.col_slot), ptr(m.owner_begin)
        wk.n_cols_dev = ptr(m.n_cols_dev)
        return wk

    def worker_step(self, mb):
        """Worker half of a step: pull -> gradient -> push of a minibatch (one launch).  `mb` is
        an index into the loaded data or an already localised minibatch."""
        m = mb if isinstance(mb, _Minibatch) else self._prepare(mb)
        self.seq += 1
        wk = self._worker_struct(m)
        grid = max(1, min(self.grid, (max(m.n_rows, m.n_cols // 8) + 7) // 8))
        check(self.lib.psb_sgd_worker_step(C.byref(wk), self.seq, grid, self.bar.data_ptr(),
                                           current_stream()))
        self._step_launches += 1
        return m

    def server_step(self) -> None:
        """Server half: apply push `seq` of every worker in rank order, acknowledge."""
        self.sv.sgd_t = math.sqrt(self.seq)
        if self.cfg.fused_server:
            check(self.lib.psb_sgd_server_update_fused(C.byref(self.sv), self.seq, self.grid_server,
                                                       self.srv_bar.data_ptr(), current_stream()))
            self._step_launches += 1
        else:
            check(self.lib.psb_sgd_server_update(C.byref(self.sv), self.seq, self.grid_server,
                                                 current_stream()))
            self._step_launches += 1 + self.world

    def check_table_load(self) -> float:
        """Open addressing with linear probing degrades sharply when a shard's table fills up
        (probe chains of thousands, each probe a peer round trip for remote workers): refuse to
        crawl.  Returns the load factor of this rank's shard."""
        code = int(self.overflow.item())
        if code == 2:
            raise RuntimeError(f"rank {self.rank}: a minibatch sends more than {self.land_cap} distinct keys to one "
                               "owner (the landing region): raise max_unique_per_minibatch")
        if code:
            raise RuntimeError(f"rank {self.rank}: a shard's key table is full (table_capacity="
                               f"{self.cfg.table_capacity}); raise AsyncSGDConfig.table_capacity")
        used = int((self.b_keys.tensor != -1).sum().item())
        load = used / self.cfg.table_capacity
        if load > 0.75:
            raise RuntimeError(f"rank {self.rank}: key table {load:.0%} full ({used} of {self.cfg.table_capacity} "
                               "slots); raise AsyncSGDConfig.table_capacity to at least twice the keys of a shard")
        return load

    def check_fault(self) -> None:
        """Raise if a device-side wait on a peer's flag timed out (the run is invalid from there)."""
        if int(self.fault[0]) != 0:
            code, idx, want, seen = (int(x) for x in self.fault.tolist())
            what = {1: "worker waiting for an owner's acknowledgement", 2: "server waiting for a worker's push"}
            raise RuntimeError(f"rank {self.rank}: device wait timed out: {what.get(code, code)} "
                               f"(rank {idx // 100}, peer {idx % 100}) wanted seq {want}, saw {seen}")

    def step(self, mb):
        """One minibatch: worker kernel, then this rank's server applies every worker's push."""
        m = self.worker_step(mb)
        self.server_step()
        self.check_fault()
        return m

    def train(self, verbose: bool = True) -> list:
        cfg = self.cfg
        t0 = time.time()
        seen = 0
        for p in range(max(1, cfg.num_data_pass)):
            for i in range(self.steps_per_pass):
                m = self.step(i if i < self.n_minibatches else self.n_minibatches)  # past the end: empty
                seen += m.n_rows
                if (i + 1) % cfg.report_every == 0 or i + 1 == self.steps_per_pass:
                    self._report(seen, time.time() - t0, m, verbose)
        return self.history

    # --------------------------------------------------------------------------------------------
    def empty_minibatch(self) -> _Minibatch:
        """What a rank pushes when it has run out of data while others have not."""
        dev = self.device
        z64 = torch.zeros(1, dtype=torch.int64, device=dev)
        return self._localize(z64, torch.empty(0, dtype=torch.int64, device=dev), None,
                              torch.empty(0, dtype=torch.float32, device=dev))

    def train_stream(self, source, steps: int, verbose: bool = False, on_step=None) -> int:
        """Streaming mode.  `source` yields host minibatches (SparseCSR whose tensors are in
        pinned memory).  Runs exactly `steps` steps (the same number on every rank; a rank whose
        source ends early pushes empty minibatches) and returns the examples consumed.  Windows of
        `cfg.stream_window` minibatches are handed to the native runner, which enqueues copy,
        localisation and both step kernels of every minibatch and returns without waiting for the
        device.  `on_step(i, m)` is called after the LAST step i of every window (`m`: that
        minibatch's device arrays, valid until the next window is submitted)."""
        it = iter(source)
        main = current_stream()
        window = max(1, min(self.cfg.stream_window, self.cfg.report_every))
        seen = done = 0
        t0 = time.time()
        ended = False
        while done < steps:
            t_a = time.perf_counter()
            hosts = []
            while not ended and len(hosts) < min(window, steps - done):
                h = next(it, None)
                if h is None:
                    ended = True
                else:
                    hosts.append(h)
            n = min(window, steps - done)
            t_b = time.perf_counter()
            self.source_s += t_b - t_a
            runner = self._stream_runner(max([h.n_rows for h in hosts] + [1]), max([h.nnz for h in hosts] + [1]),
                                         any(h.vals is not None for h in hosts))
            last = runner.submit(hosts, n, self.seq + 1, main)
            self.seq += n
            self._step_launches += n * (2 if self.cfg.fused_server else 2 + self.world)
            self._loc_launches += sum((5 if self.tail_sketch is not None else 4) for h in hosts if h.nnz)
            self.h2d_bytes += sum(8 * (h.n_rows + 1) + 4 * h.n_rows + h.nnz * (8 + (4 if h.vals is not None else 0))
                                  for h in hosts)
            self.submit_s += time.perf_counter() - t_b
            done += n
            seen += sum(h.n_rows for h in hosts)
            self.check_fault()
            self._check_overflow(runner.overflow_host)
            if on_step is not None:
                on_step(done - 1, last)
            if verbose and (done % self.cfg.report_every == 0 or done == steps):
                self._report(seen, time.time() - t0, last, verbose)
        # the load factor of the shard (a scan of the table): once per call, not per window
        self.check_table_load()
        return seen

    def _stream_runner(self, max_rows: int, max_nnz: int, with_vals: bool):
        r = self._runner
        if r is None or r.max_rows < max_rows or r.max_nnz < max_nnz or (with_vals and not r.with_vals):
            if r is not None:   # a larger minibatch than any before: drain, rebuild the ring
                max_rows, max_nnz = max(max_rows, r.max_rows), max(max_nnz, 2 * r.max_nnz)
                with_vals |= r.with_vals
                r.close()
            self._runner = r = _StreamRunner(self, max_rows, max_nnz, with_vals)
        return r

    def _check_overflow(self, flag_host) -> None:
        code = int(flag_host[0])
        if code == 2:
            raise RuntimeError(f"rank {self.rank}: a minibatch sends more than {self.land_cap} distinct keys to one "
                               "owner (the landing region): raise max_unique_per_minibatch")
        if code:
            raise RuntimeError(f"rank {self.rank}: a shard's key table is full (table_capacity="
                               f"{self.cfg.table_capacity}); raise AsyncSGDConfig.table_capacity")

    def _report(self, seen, secs, m, verbose):
        st = self.wstats.cpu().tolist()
        self.wstats.zero_()
        rows = max(st[2], 1.0)
        auc = float("nan")
        if m.n_rows:
            from ..ops.evaluation import auc as dev_auc
            auc = dev_auc(m.y, (m.xw if m.xw is not None else self._xw)[:m.n_rows])
        ss = self.sstats.cpu().tolist()
        self.sstats[1:].zero_()
        rec = dict(examples=seen, loss=st[0] / rows, accuracy=st[1] / rows, auc=auc, nnz_w=ss[0],
                   update_ratio=math.sqrt(ss[2] + 1e-20) / math.sqrt(max(ss[1], 1e-300)), sec=secs)
        self.history.append(rec)
        if verbose and self.rank == 0:
            if len(self.history) == 1:
                print(" sec  examples    loss      auc   accuracy   |w|_0  updt ratio", file=self.out)
            print("%4d  %.2e  %.3e  %.4f  %.4f  %.2e  %.2e" % (
                int(secs), seen, rec["loss"], auc, rec["accuracy"], rec["nnz_w"], rec["update_ratio"]),
                file=self.out)

    # --------------------------------------------------------------------------------------------
    def shard_model(self):
        """(keys uint64 as int64, weights) of the non-zero entries of this rank's table."""
        cap = self.cfg.table_capacity
        ok = torch.empty(cap, dtype=torch.int64, device=self.device)
        ow = torch.empty(cap, dtype=torch.float32, device=self.device)
        cnt = torch.zeros(1, dtype=torch.int64, device=self.device)
        check(self.lib.psb_kv_export_nonzero(self.b_keys.ptr, self.b_w.ptr, cap, ptr(ok), ptr(ow), ptr(cnt),
                                             current_stream()))
        n = int(cnt.item())
        order = torch.argsort(ok[:n])
        return ok[:n][order], ow[:n][order]

    def save_model(self, path_prefix: str, node_id: str | None = None) -> str:
        from ..utils.model_io import write_model_text
        k, w = self.shard_model()
        return write_model_text(f"{path_prefix}_{node_id or 'S%d' % self.rank}",
                                k.cpu().numpy().view(np.uint64), w.double().cpu().numpy())

    def close(self):
        if self._runner is not None:
            self._runner.close()
            self._runner = None
        if hasattr(self, "arena"):
            torch.cuda.synchronize(self.device)
            self.arena.close()


class _StreamRunner:
    """Owner of the device memory of the native streaming runner (csrc/cuda/sgd_stream.cu): a ring
    of buffer sets, one scratch set of the localisation and one copy stream per lane (`copy_lanes`
    minibatches are uploaded / localised concurrently), the step buffers.  The producer half of the reference's MinibatchReader (src/learner/sgd.h:120-129) with
    the H2D copy in place of the text parser's queue."""

    def __init__(self, eng: "AsyncSGDEngine", max_rows: int, max_nnz: int, with_vals: bool):
        dev, W, lib = eng.device, eng.world, eng.lib
        self.eng, self.max_rows, self.max_nnz, self.with_vals = eng, max_rows, max_nnz, with_vals
        n_lanes = max(1, min(MAX_LANES, int(os.environ.get("PSB_SGD_COPY_LANES", eng.cfg.copy_lanes))))
        self.copy_streams = [torch.cuda.Stream(dev) for _ in range(n_lanes)]
        n_sets = min(MAX_RING, max(n_lanes + 1, eng.cfg.prefetch + 1))
        i64, i32, f32 = (dict(dtype=t, device=dev) for t in (torch.int64, torch.int32, torch.float32))
        self.sets = []
        cfg = RunnerConfigStruct()
        wk = eng._worker_template(max_nnz, max_rows)   # peer views + step buffers of this capacity
        # the native side keeps these addresses: hold the tensors even if the engine replaces its own
        self.w_local, self.g_local, self.xw = eng._w_local, eng._g_local, eng._xw
        C.memmove(C.byref(cfg.worker), C.byref(wk), C.sizeof(SgdWorkerStruct))
        C.memmove(C.byref(cfg.server), C.byref(eng.sv), C.sizeof(SgdServerStruct))
        for i in range(n_sets):
            t = dict(row_ptr=torch.zeros(max_rows + 1, **i64), keys=torch.empty(max_nnz, **i64),
                     vals=torch.empty(max_nnz, **f32) if with_vals else None, y=torch.empty(max_rows, **f32),
                     cols=torch.empty(max_nnz, **i64), col=torch.empty(max_nnz, **i32),
                     col_owner=torch.empty(max_nnz, **i32), col_slot=torch.empty(max_nnz, **i64),
                     counts=torch.zeros(2, **i64), owner_begin=torch.zeros(W + 1, **i64))
            self.sets.append(t)
            for name, v in t.items():
                setattr(cfg.set[i], name, ptr(v))
        temp_bytes = int(lib.psb_sgd_localize_temp_bytes(max_nnz))
        self.scratch = [dict(sorted=torch.empty(max_nnz, **i64), uniq=torch.empty(max_nnz, **i64),
                             cnt=torch.empty(max_nnz, **i32), flag=torch.empty(max_nnz, dtype=torch.uint8, device=dev),
                             temp=torch.empty(temp_bytes, dtype=torch.uint8, device=dev)) for _ in range(n_lanes)]
        for lane, sc in enumerate(self.scratch):
            for name, v in sc.items():
                setattr(cfg.scratch[lane], name, ptr(v))
        cfg.temp_bytes = temp_bytes
        cfg.n_sets, cfg.n_lanes, cfg.fused_server = n_sets, n_lanes, int(eng.cfg.fused_server)
        cfg.max_rows, cfg.max_nnz = max_rows, max_nnz
        sk = eng.tail_sketch
        if sk is not None:
            cfg.sketch, cfg.cm_n, cfg.cm_k, cfg.freq = sk.table.data_ptr(), sk.n, sk.k, eng.cfg.tail_feature_freq
        cfg.bounds_dev = ptr(eng.bounds)
        for s in range(W):
            cfg.table_keys[s], cfg.table_cap[s] = eng.tab_keys_ptrs[s], eng.tab_caps[s]
        self.overflow_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        cfg.overflow_dev, cfg.overflow_host = ptr(eng.overflow), self.overflow_host.data_ptr()
        cfg.worker_bar, cfg.server_bar = eng.bar.data_ptr(), eng.srv_bar.data_ptr()
        cfg.grid_worker, cfg.grid_server = eng.grid, eng.grid_server
        for lane, st in enumerate(self.copy_streams):
            cfg.copy_stream[lane] = st.cuda_stream
        torch.cuda.synchronize(dev)           # the buffers above were filled on the current stream
        self.handle = lib.psb_sgd_runner_create(C.byref(cfg))
        if not self.handle:
            check(1)
        self.count = 0
        self._hosts = []                       # host arrays of the windows whose copies may still be in flight

    def submit(self, hosts: list, n: int, seq_first: int, step_stream) -> _Minibatch:
        """Enqueue `n` steps: the minibatches of `hosts`, then empty ones.  Returns the device view
        of the last one."""
        descs = (HostMinibatchStruct * n)()
        for d, h in zip(descs, hosts):
            d.row_ptr, d.keys, d.vals, d.y = h.row_ptr.data_ptr(), h.keys.data_ptr(), ptr(h.vals), h.y.data_ptr()
            d.n_rows, d.nnz = h.n_rows, h.nnz
        check(self.eng.lib.psb_sgd_runner_submit(self.handle, descs, n, seq_first, step_stream))
        copied = [torch.cuda.Event() for _ in self.copy_streams]
        for ev, st in zip(copied, self.copy_streams):
            ev.record(st)
        self._hosts.append((hosts, copied))
        while len(self._hosts) > 2:            # two windows back: its copies are long done
            for ev in self._hosts.pop(0)[1]:
                ev.synchronize()
        self.count += n
        t = self.sets[(self.count - 1) % len(self.sets)]
        rows = hosts[n - 1].n_rows if len(hosts) == n else 0
        nnz = hosts[n - 1].nnz if len(hosts) == n else 0
        return _Minibatch(t["row_ptr"][:rows + 1], t["col"][:nnz], t["vals"][:nnz] if t["vals"] is not None else None,
                          t["y"][:rows], t["cols"], t["col_owner"], t["col_slot"], t["owner_begin"], rows, nnz,
                          t["counts"][1:], self.xw)

    def timing(self) -> dict | None:
        """PSB_SGD_TIMING=1: mean device milliseconds per step of the stages (synchronises)."""
        out = (C.c_double * 7)()
        n = self.eng.lib.psb_sgd_runner_timing(self.handle, out)
        names = ("h2d", "localize", "worker", "server", "copy_stage", "step_stage", "period")
        return dict(zip(names, (round(v, 4) for v in out)), steps=n) if n else None

    def close(self):
        if self.handle:
            torch.cuda.synchronize(self.eng.device)
            self.eng.lib.psb_sgd_runner_destroy(self.handle)
            self.handle = None


def _as_i64(u: int) -> int:
    u &= 0xFFFFFFFFFFFFFFFF
    return u - (1 << 64) if u >= (1 << 63) else u
