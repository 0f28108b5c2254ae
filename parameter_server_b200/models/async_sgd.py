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
                            src/learner/sgd.h:104-157): a prefetch thread copies them from pinned
                            memory (one async copy per array) and localises them on a side stream
                            while the step kernels of earlier minibatches run; nothing is kept.
The tail-feature filter is the reference's: one CountMin sketch per worker that lives across
minibatches (insert the counts of a minibatch, keep the keys whose estimate exceeds the bound).
"""
from __future__ import annotations

import ctypes as C
import math
import os
import queue
import sys
import threading
import time
from dataclasses import dataclass

import numpy as np
import torch

from .._lib import check, cuda_lib, current_stream, ptr
from ..ops import localizer as L
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
                ("depth", C.c_int32), ("loss", C.c_int32), ("quant", C.c_int32)]


class SgdServerStruct(C.Structure):
    _fields_ = [("w", P), ("z", P), ("n", P), ("cap", C.c_int64), ("land_slot", P), ("land_val", P),
                ("land_q", P), ("land_scale", P), ("land_cnt", P), ("land_flag", P),
                ("ack_out", P * MAX_PEERS), ("land_cap", C.c_int64), ("stats", P),
                ("rank", C.c_int32), ("n_workers", C.c_int32), ("depth", C.c_int32),
                ("algo", C.c_int32), ("quant", C.c_int32), ("alpha", C.c_float), ("beta", C.c_float),
                ("l1", C.c_float), ("l2", C.c_float), ("sgd_t", C.c_float), ("lr_constant", C.c_int32)]


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
    prefetch: int = 3                # streaming mode: localised minibatches kept ahead of the step
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
    n_cols: int


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
        self.launches = 0
        self.h2d_bytes = 0
        self.prefetch_busy_s = 0.0   # streaming: wall time the prefetch thread spent localising
        self.feed_wait_s = 0.0       # streaming: wall time the step loop waited for the prefetcher
        self.data = None
        self.n_rows = 0
        # PSB_SGD_PROFILE=1: wall time per localisation stage (synchronising after each: slower)
        self._stage_prof = {} if os.environ.get("PSB_SGD_PROFILE") else None

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
        """K10 + K13 + K11 + K15 + K8a on the current stream: unique keys and counts, tail filter,
        remap to local columns, owner segments, slots in the owners' tables."""
        cfg, dev = self.cfg, self.device
        prof = self._stage_prof

        def lap(name, _t=[time.perf_counter()]):
            if prof is None:
                return
            torch.cuda.current_stream(dev).synchronize()
            now = time.perf_counter()
            prof[name] = prof.get(name, 0.0) + now - _t[0]
            _t[0] = now
        lap("h2d_wait")
        nnz, n_rows = keys_nnz.numel(), row_ptr.numel() - 1
        uniq, cnt = L.unique_count(keys_nnz) if nnz else (keys_nnz, None)
        lap("sort_unique")
        if self.tail_sketch is not None and nnz:
            self.tail_sketch.insert(uniq, cnt.clamp(max=255))
            uniq = uniq[self.tail_sketch.query_gt(uniq, cfg.tail_feature_freq)].contiguous()
        n_cols = uniq.numel()
        lap("tail_filter")
        idx1 = L.lookup_sorted(uniq, keys_nnz) if nnz else torch.empty(0, dtype=torch.int32, device=dev)
        rp, col, val = L.compact_csr(row_ptr, idx1, vals)
        lap("remap")
        # owners: the sorted unique keys split at the shard boundaries
        W = self.world
        if W > 1 and n_cols:
            cut = L.lower_bound(uniq, self.bounds).tolist()
        else:
            cut = [n_cols] * (W - 1)
        ob = [0] + cut + [n_cols]
        owner_begin = torch.tensor(ob, dtype=torch.int64, device=dev)
        col_owner = torch.repeat_interleave(torch.arange(W, dtype=torch.int32, device=dev),
                                            owner_begin[1:] - owner_begin[:-1]) if n_cols else \
            torch.zeros(1, dtype=torch.int32, device=dev)
        col_slot = torch.full((max(n_cols, 1),), -1, dtype=torch.int64, device=dev)
        lap("owners")
        if n_cols:
            assert max(ob[s + 1] - ob[s] for s in range(W)) <= self.land_cap, \
                "landing region too small: raise max_unique_per_minibatch"
            check(self.lib.psb_kv_lookup_insert(ptr(uniq), n_cols, ptr(col_owner), self.tab_keys_ptrs,
                                                self.tab_caps, W, ptr(col_slot), ptr(self.overflow),
                                                current_stream()))
        lap("table_slots")
        self._new_minibatches = getattr(self, "_new_minibatches", 0) + 1
        if self._new_minibatches % 64 == 0:
            self.check_table_load()
        lap("load_check")
        return _Minibatch(rp, col, val, y, uniq, col_owner, col_slot, owner_begin, n_rows, n_cols)

    def _worker_struct(self, m: _Minibatch):
        cfg, A, W, D = self.cfg, self.arena, self.world, self.depth
        dev = self.device
        nblk = self.land_cap // 128 + 1
        self._w_local = torch.empty(max(m.n_cols, 1), dtype=torch.float32, device=dev)
        self._g_local = torch.empty(max(m.n_cols, 1), dtype=torch.float32, device=dev)
        self._xw = torch.empty(max(m.n_rows, 1), dtype=torch.float32, device=dev)
        wk = SgdWorkerStruct()
        wk.row_ptr, wk.col, wk.val, wk.y = ptr(m.row_ptr), ptr(m.col), ptr(m.val), ptr(m.y)
        wk.n_rows, wk.n_cols = m.n_rows, m.n_cols
        wk.col_owner, wk.col_slot, wk.owner_begin = ptr(m.col_owner), ptr(m.col_slot), ptr(m.owner_begin)
        wk.w_local, wk.g_local, wk.xw = ptr(self._w_local), ptr(self._g_local), ptr(self._xw)
        wk.stats = self.wstats.data_ptr()
        r = self.rank
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
        self.launches += 1
        return m

    def server_step(self) -> None:
        """Server half: apply push `seq` of every worker in rank order, acknowledge."""
        self.sv.sgd_t = math.sqrt(self.seq)
        if self.cfg.fused_server:
            check(self.lib.psb_sgd_server_update_fused(C.byref(self.sv), self.seq, self.grid_server,
                                                       self.srv_bar.data_ptr(), current_stream()))
            self.launches += 1
        else:
            check(self.lib.psb_sgd_server_update(C.byref(self.sv), self.seq, self.grid_server,
                                                 current_stream()))
            self.launches += 1 + self.world

    def check_table_load(self) -> float:
        """Open addressing with linear probing degrades sharply when a shard's table fills up
        (probe chains of thousands, each probe a peer round trip for remote workers): refuse to
        crawl.  Returns the load factor of this rank's shard."""
        if int(self.overflow.item()):
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
        pinned memory); a prefetch thread uploads and localises them `cfg.prefetch` ahead on its
        own stream.  Runs exactly `steps` steps (the same number on every rank; a rank whose
        source ends early pushes empty minibatches).  Returns the examples consumed."""
        feeder = _Prefetcher(self, source, self.cfg.prefetch)
        feeder.start()
        main = torch.cuda.current_stream(self.device)
        seen = 0
        t0 = time.time()
        try:
            for i in range(steps):
                t_q = time.perf_counter()
                item = feeder.get()
                self.feed_wait_s += time.perf_counter() - t_q
                if item is None:
                    m = self.empty_minibatch()
                else:
                    m, ready = item
                    main.wait_event(ready)
                    for t in (m.row_ptr, m.col, m.val, m.y, m.keys, m.col_owner, m.col_slot, m.owner_begin):
                        if t is not None:
                            t.record_stream(main)  # allocated on the prefetch stream, read here
                self.step(m)
                seen += m.n_rows
                if on_step is not None:
                    on_step(i, m)
                if verbose and ((i + 1) % self.cfg.report_every == 0 or i + 1 == steps):
                    self._report(seen, time.time() - t0, m, verbose)
        finally:
            feeder.stop()
        return seen

    def _report(self, seen, secs, m, verbose):
        st = self.wstats.cpu().tolist()
        self.wstats.zero_()
        rows = max(st[2], 1.0)
        auc = float("nan")
        if m.n_rows:
            from ..ops.evaluation import auc as dev_auc
            auc = dev_auc(m.y, self._xw[:m.n_rows])
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
        if hasattr(self, "arena"):
            torch.cuda.synchronize(self.device)
            self.arena.close()


class _Prefetcher(threading.Thread):
    """Host minibatch -> HBM -> localised, ahead of the step loop (the producer half of the
    reference's MinibatchReader, src/learner/sgd.h:120-129, with the H2D copy in place of the text
    parser's queue).  One cudaMemcpyAsync per array from pinned memory on a dedicated stream; the
    host syncs of the localisation (sizes of the unique-key lists) block only this thread."""

    def __init__(self, engine, source, depth: int):
        super().__init__(daemon=True)
        self.eng, self.source = engine, iter(source)
        self.q = queue.Queue(maxsize=max(1, depth))
        self.stream = torch.cuda.Stream(engine.device)
        self._halt = threading.Event()
        self._ended = False
        self.error = None

    def run(self):
        eng, dev = self.eng, self.eng.device
        try:
            torch.cuda.set_device(dev)
            with torch.cuda.stream(self.stream):
                for host in self.source:
                    if self._halt.is_set():
                        return
                    up = lambda t: None if t is None else t.to(dev, non_blocking=True)  # noqa: E731
                    rp, keys, vals, y = up(host.row_ptr), up(host.keys), up(host.vals), up(host.y)
                    eng.h2d_bytes += sum(t.numel() * t.element_size() for t in (rp, keys, vals, y) if t is not None)
                    t_loc = time.perf_counter()
                    m = eng._localize(rp, keys, vals, y)
                    eng.prefetch_busy_s += time.perf_counter() - t_loc
                    ready = torch.cuda.Event()
                    ready.record(self.stream)
                    while not self._halt.is_set():
                        try:
                            self.q.put((m, ready), timeout=0.1)
                            break
                        except queue.Full:
                            continue
        except BaseException as e:  # noqa: BLE001 - surfaced by get()
            self.error = e
        finally:
            while not self._halt.is_set():
                try:
                    self.q.put(None, timeout=0.1)
                    break
                except queue.Full:
                    continue

    def get(self):
        """The next localised minibatch, or None once the source is exhausted (and from then on)."""
        if self._ended:
            return None
        item = self.q.get()
        if self.error is not None:
            raise self.error
        if item is None:
            self._ended = True
        return item

    def stop(self):
        self._halt.set()
        while True:
            try:
                self.q.get_nowait()
            except queue.Empty:
                break
        self.join(timeout=10)


def _as_i64(u: int) -> int:
    u &= 0xFFFFFFFFFFFFFFFF
    return u - (1 << 64) if u >= (1 << 63) else u
