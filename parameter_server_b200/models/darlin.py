"""Darlin on B200: L1 logistic regression by (bounded-delay) block coordinate descent.

Every rank (one process per GPU) is a worker *and* a server shard.  The scheduler logic of
the reference (DarlinScheduler::run, src/app/linear_method/darlin.h:32-132 and
BCDScheduler, src/learner/bcd.h:26-191) runs replicated and deterministically on every
rank; the per-block push / update / pull chain runs entirely on the device:

    grad_push -> [release flags over NVLink] -> server_update -> [flags] -> pull_update

A whole pass over the feature blocks is captured once into a CUDA graph and replayed with a
new block order / KKT threshold (small H2D copy) every pass.  Bounded delay tau maps to
merging tau+1 consecutive blocks of the order into one step: their gradients are computed
against the same (stale by <= tau blocks) margins, which is one admissible schedule of the
reference's wait_time protocol (executor.cc:168-180) and divides the number of dependent
cross-GPU synchronisations by tau+1.
"""
from __future__ import annotations

import ctypes as C
import math
import sys
import time
from dataclasses import dataclass, field

import numpy as np
import torch

from ..ops import darlin as K
from ..ops import localizer as L
from ..parallel.comm import Comm, SingleComm
from ..utils.range import Range
from ..utils.synthetic import SparseCSR


@dataclass
class DarlinConfig:
    """Subset of LM::Config + BCDConfig that drives Darlin (linear.proto, bcd.proto)."""
    lambda_l1: float = 1.0            # penalty.lambda(0)
    eta: float = 1.0                  # learning_rate.alpha
    max_pass_of_data: int = 10
    epsilon: float = 1e-4
    feature_block_ratio: float = 4.0
    random_feature_block_order: bool = True
    max_block_delay: int = 0          # tau
    tail_feature_freq: int = 0
    countmin_k: int = 2               # bcd.proto:30
    countmin_n_ratio: float = 2.0     # bcd.proto:32
    tail_filter: str = "exact"        # "exact": true global counts (device sort + RLE);
    #                                   "countmin": the reference's sketch (may keep a few rare
    #                                   features because of hash collisions), same hash as the host
    prior_fea_group: list = field(default_factory=list)
    num_iter_for_prior_fea_group: int = 5
    delta_init_value: float = 1.0
    delta_max_value: float = 5.0
    kkt_filter_threshold_ratio: float = 10.0
    # B200-specific knobs (no counterpart in the reference)
    seg_len: int = 128                # nnz per warp work segment
    shard_by: str = "data_range"      # "data_range": even split of [min_key,max_key);
    #                                   "all": Range::all() like postmaster.cc:17-31
    revive_on_reset: bool = False
    seed: int = 0
    use_graph: bool = True
    fused: bool = True                # one persistent cooperative kernel per pass
    kernel_flags: int = 0             # psb_darlin_ctrl.flags
    push_format: str = "fp64"         # multi-rank pass: "fp32" rounds the pushed (g, u) (lossy, half the bytes)
    collectives: str = "peer"         # "peer": kernels store / load peer memory directly;
    #                                   "nccl": the comparison variant (split kernels with
    #                                   all_to_all + all_gather between them, models/darlin_nccl.py)
    fused_grid: int = 0               # 0 = all co-resident CTAs
    profile_phases: bool = False      # accumulate per-phase %globaltimer time (debug)

    @staticmethod
    def from_proto_text(conf_text: str) -> "DarlinConfig":
        """From the text of an AppConfig (`linear_method { ... darlin { ... } }`, linear.proto /
        bcd.proto incl. the `[PS.LM.*]` extensions); fields that are absent keep the proto defaults."""
        from ..runtime.host import textproto_get
        base = "linear_method."
        g = lambda k: textproto_get(conf_text, base + k)  # noqa: E731
        c = DarlinConfig()
        if g("penalty.lambda[0]"):
            c.lambda_l1 = float(g("penalty.lambda[0]"))
        if g("learning_rate.alpha"):
            c.eta = float(g("learning_rate.alpha"))
        for name in ("max_pass_of_data", "epsilon", "feature_block_ratio", "max_block_delay", "tail_feature_freq",
                     "num_iter_for_prior_fea_group", "countmin_k", "countmin_n_ratio"):
            v = g("darlin." + name)
            if v:
                setattr(c, name, type(getattr(c, name))(float(v)))
        v = g("darlin.random_feature_block_order")
        if v:
            c.random_feature_block_order = v == "true"
        for ext in ("delta_init_value", "delta_max_value", "kkt_filter_threshold_ratio"):
            v = g(f"darlin.[PS.LM.{ext}]")
            if v:
                setattr(c, ext, float(v))
        i = 0
        while g(f"darlin.prior_fea_group[{i}]"):
            c.prior_fea_group.append(int(g(f"darlin.prior_fea_group[{i}]")))
            i += 1
        return c


@dataclass
class Progress:
    objective: float = 0.0
    relative_obj: float = 1.0
    nnz_w: int = 0
    violation: float = 0.0
    nnz_active_set: int = 0
    total_time: float = 0.0
    busy_time: float = 0.0
    device_ms: float = 0.0   # H2D + kernels + D2H of the pass (CUDA events)
    kernel_ms: float = 0.0   # kernels of the pass only


class DarlinEngine:
    """Worker + server shard of one rank.  See the module docstring."""

    def __init__(self, config: DarlinConfig, comm: Comm | None = None, device=None,
                 out=sys.stderr):
        self.cfg = config
        self.comm = comm if comm is not None else SingleComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        self.device = torch.device(device if device is not None else
                                   f"cuda:{torch.cuda.current_device()}")
        self.out = out
        self.progress: list[Progress] = []
        self._graphs = {}
        self._pass_index = 0
        self._seq = 0
        self.launches_per_pass = 0
        self.h2d_bytes_per_pass = 0
        self.d2h_bytes_per_pass = 0

    # ----------------------------------------------------------------------------------------
    # LOAD_DATA + PREPROCESS_DATA  (ref bcd.h:32-76, 324-465; localizer.h)
    # ----------------------------------------------------------------------------------------
    def load(self, groups: dict | SparseCSR) -> None:
        if isinstance(groups, SparseCSR):
            groups = {1: groups}
        cfg, dev, W = self.cfg, self.device, self.world
        grp_ids = sorted(groups)
        first = groups[grp_ids[0]]
        self.y = first.y.to(dev).float().contiguous()
        self.n_rows = first.n_rows
        tot = self.comm.allreduce_sum([float(self.n_rows)], dev)
        self.num_ex_global = int(tot[0])

        # K10: sorted unique keys + counts of this worker, per group
        local = {}
        for gid in grp_ids:
            local[gid] = L.unique_count(groups[gid].to(dev).keys)
        # server key ranges (ref Postmaster::partitionServerKeyRange, postmaster.cc:17-31)
        if cfg.shard_by == "all":
            whole = Range.all()
        else:
            ends = [(int(u[0].item()) & _M64, int(u[-1].item()) & _M64) for u, _ in local.values() if u.numel()]
            ends = [e for part in self.comm.gather_obj(ends) for e in part]
            whole = Range(min(e[0] for e in ends), max(e[1] for e in ends) + 1)
        self.server_ranges = [whole.even_divide(W, s) for s in range(W)]
        cuts = torch.tensor([_as_i64(r.begin) for r in self.server_ranges[1:]], dtype=torch.int64, device=dev)

        dicts, col_base, info = {}, {}, {}
        ncols = 0
        for gid in grp_ids:
            d = groups[gid].to(dev)
            uniq, cnt = local[gid]
            if W > 1:
                # C5 / C6: every worker sends the keys (and counts) of each server's range to that
                # server only; the owner forms the union, sums the counts, applies the tail
                # filter, and the shard dictionaries are gathered into the global one.  Nobody
                # ever holds every worker's key list (ref bcd.h:353-397, shared_parameter.h:121-146).
                cut = [0] + (L.lower_bound(uniq, cuts).tolist() if uniq.numel() else [0] * (W - 1)) + [uniq.numel()]
                got_k = self.comm.alltoall_var([uniq[cut[s]:cut[s + 1]].contiguous() for s in range(W)])
                got_c = self.comm.alltoall_var([cnt[cut[s]:cut[s + 1]].contiguous() for s in range(W)])
                mine = L.unique_count(torch.cat(got_k))[0]
            else:
                got_k, got_c, mine = [uniq], [cnt], uniq
            if cfg.tail_feature_freq > 0 and cfg.tail_filter == "countmin":
                # the reference's sketch on the owner: insert every worker's (key, count), then query
                from ..ops.evaluation import CountMin
                n = int(self.comm.allreduce_sum([float(max(p.numel() for p in got_k))], dev)[0] * cfg.countmin_n_ratio)
                cm = CountMin(max(n, 16), cfg.countmin_k, dev)
                for p, c in zip(got_k, got_c):
                    if p.numel():
                        cm.insert(p, c.clamp(max=255))
                if mine.numel():
                    mine = mine[cm.query_gt(mine, cfg.tail_feature_freq)].contiguous()
            elif cfg.tail_feature_freq > 0:
                total = torch.zeros(mine.numel(), dtype=torch.int64, device=dev)
                for p, c in zip(got_k, got_c):
                    if p.numel():
                        total.index_add_(0, L.lookup_sorted(mine, p).long() - 1, c.long())
                mine = mine[total > cfg.tail_feature_freq].contiguous()
            dictionary = torch.cat([t.to(dev) for t in self.comm.gather_var(mine)]) if W > 1 else mine
            nnz_ele, nnz_ex = float(d.nnz), float((d.row_ptr[1:] > d.row_ptr[:-1]).sum().item())
            s = self.comm.allreduce_sum([nnz_ele, nnz_ex], dev)
            info[gid] = dict(nnz_ele=s[0], nnz_ex=s[1])
            dicts[gid] = dictionary
            col_base[gid] = ncols
            ncols += dictionary.numel()
        self.grp_ids, self.dicts, self.col_base, self.n_cols = grp_ids, dicts, col_base, ncols
        self.keys_all = torch.cat([dicts[g] for g in grp_ids]) if grp_ids else None

        # K11 + K12 per group, then concatenate the CSC pieces
        cps, ris, svs = [], [], []
        nnz_off = 0
        binary = all(groups[g].vals is None for g in grp_ids)
        for gid in grp_ids:
            d = groups[gid].to(dev)
            idx1 = L.lookup_sorted(dicts[gid], d.keys)
            vals = d.vals if d.vals is not None else (
                None if binary else torch.ones(d.nnz, dtype=torch.float32, device=dev))
            rp, col, val = L.compact_csr(d.row_ptr, idx1, vals)
            cp, ri, sv = L.csr_to_csc(rp, col, val, dicts[gid].numel(), self.y)
            cps.append(cp[:-1] + nnz_off if gid != grp_ids[-1] else cp + nnz_off)
            ris.append(ri)
            if sv is not None:
                svs.append(sv)
            nnz_off += ri.numel()
        self.col_ptr = torch.cat(cps).contiguous()
        self.row_idx = torch.cat(ris).contiguous()
        self.sval = torch.cat(svs).contiguous() if not binary else None
        self.nnz = nnz_off
        self.col_seg, self.seg_col, self.seg_beg, self.seg_meta = L.build_segments(self.col_ptr, cfg.seg_len)

        self._divide_blocks(info)
        self._partition_servers()
        self._allocate_state()

    def _key_at(self, gid: int, pos: int) -> int:
        return int(self.dicts[gid][pos].item()) & 0xFFFFFFFFFFFFFFFF

    def _divide_blocks(self, info: dict) -> None:
        """ref BCDScheduler::divideFeatureBlocks (bcd.h:78-122)."""
        cfg, dev = self.cfg, self.device
        self.blocks = []  # (gid, Range key range, col_lo, col_hi)
        for gid in self.grp_ids:
            n_keys = self.dicts[gid].numel()
            if n_keys == 0:
                continue
            kmin, kmax = self._key_at(gid, 0), self._key_at(gid, n_keys - 1) + 1
            npr = info[gid]["nnz_ele"] / max(info[gid]["nnz_ex"], 1.0)
            n = 1
            if npr > 1 + 1e-6:
                n = max(int(math.ceil(npr * cfg.feature_block_ratio)), 1)
            rngs = [Range(kmin, kmax).even_divide(n, i) for i in range(n)]
            rngs = [r for r in rngs if not r.empty()]
            probes = torch.tensor([_as_i64(r.begin) for r in rngs] + [_as_i64(rngs[-1].end)],
                                  dtype=torch.int64, device=dev)
            pos = L.lower_bound(self.dicts[gid], probes).tolist()
            for i, r in enumerate(rngs):
                self.blocks.append((gid, r, self.col_base[gid] + pos[i], self.col_base[gid] + pos[i + 1]))
        self.blk_order = list(range(len(self.blocks)))
        rng = np.random.RandomState(cfg.seed + 17)
        self.prior_blk_order = []
        for gid in cfg.prior_fea_group:
            tmp = [k for k, b in enumerate(self.blocks) if b[0] == gid]
            if not tmp:
                continue
            for _ in range(cfg.num_iter_for_prior_fea_group):
                if cfg.random_feature_block_order:
                    rng.shuffle(tmp)
                self.prior_blk_order += list(tmp)
        self._order_rng = np.random.RandomState(cfg.seed + 23)

    def _partition_servers(self) -> None:
        """ref Postmaster::partitionServerKeyRange (postmaster.cc:17-31) + sliceKeyOrderedMsg."""
        cfg, dev, S = self.cfg, self.device, self.world
        bounds = [r.begin for r in self.server_ranges] + [self.server_ranges[-1].end]
        probes = torch.tensor([_as_i64(b) for b in bounds], dtype=torch.int64, device=dev)
        # slice[g][s] = [lo, hi) positions of server s inside group g's dictionary
        self.slices = {}
        shard_len = [0] * S
        self.shard_base = {}
        for gid in self.grp_ids:
            pos = L.lower_bound(self.dicts[gid], probes).tolist()
            if cfg.shard_by == "all":
                pos[-1] = self.dicts[gid].numel()  # Range::all() excludes 2^64-1 only formally
            self.slices[gid] = [(pos[s], pos[s + 1]) for s in range(S)]
            for s in range(S):
                self.shard_base[(gid, s)] = shard_len[s]
                shard_len[s] += pos[s + 1] - pos[s]
        self.shard_len = shard_len
        col_owner = torch.empty(self.n_cols, dtype=torch.int32, device=dev)
        col_off = torch.empty(self.n_cols, dtype=torch.int64, device=dev)
        for gid in self.grp_ids:
            cb = self.col_base[gid]
            for s, (lo, hi) in enumerate(self.slices[gid]):
                if hi > lo:
                    col_owner[cb + lo:cb + hi] = s
                    col_off[cb + lo:cb + hi] = torch.arange(hi - lo, device=dev) + self.shard_base[(gid, s)]
        self.col_owner, self.col_off = col_owner, col_off
        nb = len(self.blocks)
        # block x server intersections, vectorised (ref sliceKeyOrderedMsg, message.h:120-159)
        blk_col = np.zeros((max(nb, 1), 2), dtype=np.int64)
        blk_rng = np.zeros((max(nb, 1), 2), dtype=np.int64)
        blk_mask = np.zeros(max(nb, 1), dtype=np.int64)
        if nb:
            blk_col[:nb] = np.array([(b[2], b[3]) for b in self.blocks], dtype=np.int64)
            gids = np.array([b[0] for b in self.blocks])
            for gid in self.grp_ids:
                sel = np.nonzero(gids == gid)[0]
                if sel.size == 0:
                    continue
                cb = self.col_base[gid]
                plo, phi = blk_col[sel, 0] - cb, blk_col[sel, 1] - cb
                lo_s = np.array([lo for lo, _ in self.slices[gid]], dtype=np.int64)
                hi_s = np.array([hi for _, hi in self.slices[gid]], dtype=np.int64)
                a = np.maximum(plo[:, None], lo_s[None, :])
                z = np.minimum(phi[:, None], hi_s[None, :])
                inter = z > a
                blk_mask[sel] = (inter.astype(np.int64) << np.arange(S)[None, :]).sum(1)
                me = inter[:, self.rank]
                base = self.shard_base[(gid, self.rank)] - lo_s[self.rank]
                blk_rng[sel[me], 0] = base + a[me, self.rank]
                blk_rng[sel[me], 1] = base + z[me, self.rank]
                # the same for every server (used by the NCCL comparison variant)
                bases = np.array([self.shard_base[(gid, t)] for t in range(S)], dtype=np.int64) - lo_s
                rng_all = np.stack([np.where(inter, bases[None, :] + a, 0),
                                    np.where(inter, bases[None, :] + z, 0)], axis=2)
                self.blk_srv_rng = getattr(self, "blk_srv_rng", None)
                if self.blk_srv_rng is None or self.blk_srv_rng.shape[0] != nb:
                    self.blk_srv_rng = np.zeros((nb, S, 2), dtype=np.int64)
                self.blk_srv_rng[sel] = rng_all
        max_blk_keys = int(max(1, (blk_rng[:, 1] - blk_rng[:, 0]).max()))
        cs = self.col_seg.cpu().numpy()
        blk_tab = np.zeros((max(nb, 1), 6), dtype=np.int64)
        blk_tab[:, 0] = cs[blk_col[:, 0]]
        blk_tab[:, 1] = cs[blk_col[:, 1]] - cs[blk_col[:, 0]]
        blk_tab[:, 2] = blk_rng[:, 0]
        blk_tab[:, 3] = blk_rng[:, 1] - blk_rng[:, 0]
        blk_tab[:, 4] = blk_mask
        self.max_blk_segs = int(max(1, blk_tab[:, 1].max()))
        blk_tab = torch.from_numpy(blk_tab.reshape(-1))
        blk_col = torch.from_numpy(blk_col.reshape(-1))
        blk_rng = torch.from_numpy(blk_rng.reshape(-1))
        blk_mask = torch.from_numpy(blk_mask.astype(np.int32))
        self.blk_tab = blk_tab.to(dev)
        self.blk_col, self.blk_rng = blk_col.to(dev), blk_rng.to(dev)
        self.blk_srv_mask = blk_mask.to(dev)
        self.max_blk_keys = max_blk_keys

    def _allocate_state(self) -> None:
        cfg, dev, W = self.cfg, self.device, self.world
        cap = max(max(self.shard_len), 1)
        self.land_cap = cap
        self.bps = max(1, min(cfg.max_block_delay + 1, 32))
        ncols = max(self.n_cols, 1)
        nbytes = (2 * W * cap + 2 * cap) * 8 + 2 * W * ncols * 32 + 8 * W * 8 + W * 48 + 64 * 256
        self.arena = self.comm.make_arena(nbytes, dev)
        A = self.arena
        self.b_land_g = A.alloc("land_g", W * cap, torch.float64)
        self.b_land_u = A.alloc("land_u", W * cap, torch.float64)
        self.b_w = A.alloc("w", cap, torch.float64)
        self.b_dw = A.alloc("dw", cap, torch.float64)
        # fused multi-rank pass: tagged (g, u) entries [2 parities][W workers][n_cols], written by
        # the workers of every rank (psb200.h), and the flow-control flags
        self.b_land2 = A.alloc("land2", 2 * W * ncols * 4, torch.float64)
        self.b_done = A.alloc("done", W, torch.int64)
        self.b_push_flag = A.alloc("push_flag", W, torch.int64)
        self.b_ver_flag = A.alloc("ver_flag", W, torch.int64)
        self.b_bar_flag = A.alloc("bar_flag", W, torch.int64)
        self.b_progress = A.alloc("progress", W * 6, torch.int64)
        n_keys = self.shard_len[self.rank]
        self.n_keys = n_keys
        f64 = dict(dtype=torch.float64, device=dev)
        self.srv_delta = torch.full((max(n_keys, 1),), cfg.delta_init_value, **f64)
        self.srv_active = _full_bitmap(n_keys, dev)
        self.violation = torch.zeros(1, **f64)
        self.z = torch.zeros(max(self.n_rows, 1), **f64)       # init_w = ZERO  => dual = 1
        self.w_cache = torch.zeros(max(self.n_cols, 1), **f64)
        self.wk_delta = torch.full((max(self.n_cols, 1),), cfg.delta_init_value, **f64)
        self.wk_active = _full_bitmap(self.n_cols, dev)
        self.part = torch.zeros(2 * self.bps * max(self.max_blk_segs, 1), **f64)
        self.stash = torch.zeros(2 * self.bps * max(self.max_blk_segs, 1), **f64)
        self.acc = torch.zeros(4 * max(self.n_cols, 1), **f64)   # one-rank pass: (g, u) accumulators, 2 parities
        self.col_cnt = torch.zeros(max(self.n_cols, 1), dtype=torch.int32, device=dev)
        self.ctrs = torch.zeros(16, dtype=torch.int32, device=dev)
        sm = torch.cuda.get_device_properties(dev).multi_processor_count
        self.grid_worker = int(max(1, min(math.ceil(self.bps * self.max_blk_segs / 8), sm * 8)))
        self.grid_server = int(max(1, min(math.ceil(self.bps * self.max_blk_keys / 256), sm)))
        self.grid_eval = int(sm * 2)
        self.grid_fused = K.fused_max_grid(self.sval is None) if cfg.fused else 0
        if cfg.fused_grid > 0:
            self.grid_fused = min(self.grid_fused, cfg.fused_grid)
        self.bar = torch.zeros(4, dtype=torch.int32, device=dev)
        self.prof = torch.zeros(8 + 4 * (len(self.blocks) + len(self.prior_blk_order) + 8) + 64, dtype=torch.int64, device=dev)
        self.eval_scratch = torch.zeros(
            K.evaluate_scratch_bytes(max(self.grid_eval, 4 * sm)) // 8 + 2, **f64)
        # host-pinned staging for the per-pass inputs / result
        max_steps = 2 * (len(self.blocks) + len(self.prior_blk_order)) + 8  # room for repeated blocks
        self.order_host = torch.full((max_steps * self.bps,), -1, dtype=torch.int32).pin_memory()
        self.order_dev = torch.full((max_steps * self.bps,), -1, dtype=torch.int32, device=dev)
        self.ctrl_host = torch.zeros(C.sizeof(K.DarlinCtrl), dtype=torch.uint8).pin_memory()
        self.ctrl_dev = torch.zeros(C.sizeof(K.DarlinCtrl), dtype=torch.uint8, device=dev)
        self.prog_host = torch.zeros(W * 6, dtype=torch.int64).pin_memory()
        self.epoch_dev = torch.zeros(1, dtype=torch.int64, device=dev)

        wk = K.DarlinWorker()
        wk.col_ptr, wk.row_idx = self.col_ptr.data_ptr(), self.row_idx.data_ptr()
        wk.sval = self.sval.data_ptr() if self.sval is not None else None
        wk.y, wk.z, wk.w = self.y.data_ptr(), self.z.data_ptr(), self.w_cache.data_ptr()
        wk.delta, wk.active = self.wk_delta.data_ptr(), self.wk_active.data_ptr()
        wk.col_owner, wk.col_off = self.col_owner.data_ptr(), self.col_off.data_ptr()
        wk.seg_col, wk.seg_beg, wk.col_seg = (self.seg_col.data_ptr(), self.seg_beg.data_ptr(),
                                              self.col_seg.data_ptr())
        wk.blk_col, wk.blk_srv_mask = self.blk_col.data_ptr(), self.blk_srv_mask.data_ptr()
        wk.part, wk.col_cnt, wk.done_ctr = (self.part.data_ptr(), self.col_cnt.data_ptr(),
                                            self.ctrs.data_ptr())
        wk.n_rows, wk.n_cols, wk.seg_len = self.n_rows, self.n_cols, cfg.seg_len
        wk.rank, wk.n_workers, wk.n_servers = self.rank, W, W
        for s in range(W):
            wk.land_g[s] = A.peer_ptr(s, self.b_land_g)
            wk.land_u[s] = A.peer_ptr(s, self.b_land_u)
            wk.land_cap[s] = cap
            wk.push_flag[s] = A.peer_ptr(s, self.b_push_flag)
            wk.srv_w[s] = A.peer_ptr(s, self.b_w)
            wk.srv_dw[s] = A.peer_ptr(s, self.b_dw)
        wk.ver_flag = self.b_ver_flag.ptr
        wk.seg_meta = self.seg_meta.data_ptr()
        self.col_has = self._column_holders()
        wk.col_has, wk.stash = self.col_has.data_ptr(), self.stash.data_ptr()
        wk.done_local = self.b_done.ptr
        wk.push_fp32 = int(cfg.push_format == "fp32")
        wk.acc = self.acc.data_ptr()
        for s in range(W):
            wk.land2[s] = A.peer_ptr(s, self.b_land2)
            wk.done_peer[s] = A.peer_ptr(s, self.b_done, self.rank)
        self.wk = wk

        sv = K.DarlinServer()
        sv.w, sv.dw = self.b_w.ptr, self.b_dw.ptr
        sv.delta, sv.active = self.srv_delta.data_ptr(), self.srv_active.data_ptr()
        sv.land_g, sv.land_u, sv.land_cap = self.b_land_g.ptr, self.b_land_u.ptr, cap
        sv.push_flag, sv.blk_rng = self.b_push_flag.ptr, self.blk_rng.data_ptr()
        sv.violation, sv.done_ctr = self.violation.data_ptr(), self.ctrs.data_ptr() + 16
        sv.n_keys, sv.rank, sv.n_workers, sv.n_servers = n_keys, self.rank, W, W
        for r in range(W):
            sv.ver_flag[r] = A.peer_ptr(r, self.b_ver_flag, self.rank)
        # the replicated multi-rank pass and the split kernels keep the workers' copies of the
        # filter marks differently: an engine runs one or the other
        self.replicated = bool(cfg.fused and W > 1 and self.grid_fused > 0)
        self._passes_run = 0
        self.sv = sv
        self.prog_peer = torch.tensor([A.peer_ptr(r, self.b_progress, 6 * self.rank)
                                       for r in range(W)], dtype=torch.int64, device=dev)
        self.bar_peers = (C.c_void_p * W)(*[A.peer_ptr(r, self.b_bar_flag) for r in range(W)])
        self._warm_kernels()
        torch.cuda.synchronize(dev)
        self.comm.barrier()

    def _column_holders(self) -> torch.Tensor:
        """col_has[j] bit r: worker r holds nonzeros of column j (it pushes the column whenever
        the column is active).  Set-up time exchange: one byte per (worker, column)."""
        present = (self.col_ptr[1:] > self.col_ptr[:-1]).to(torch.uint8)
        has = torch.zeros(max(self.n_cols, 1), dtype=torch.int32, device=self.device)
        for r, p in enumerate(self.comm.gather_var(present)):
            has[:p.numel()] |= p.to(self.device).to(torch.int32) << r
        return has

    def _warm_kernels(self) -> None:
        """Load every kernel once with an empty step so that graph capture never triggers a
        lazy module load; an all -1 order touches no state besides violation/progress."""
        ctrl = K.DarlinCtrl(1e20, self.cfg.lambda_l1, self.cfg.eta, self.cfg.delta_max_value, 0, 0, 0, 0, 0, 0, 0)
        self.ctrl_dev.copy_(torch.frombuffer(bytearray(bytes(ctrl)), dtype=torch.uint8))
        cp, op = self.ctrl_dev.data_ptr(), self.order_dev.data_ptr()
        K.pass_begin(self.wk, self.sv, cp)
        K.grad_push(self.wk, cp, op, 0, self.bps, self.grid_worker)
        K.server_update(self.sv, cp, op, 0, self.bps, self.grid_server)
        K.pull_update(self.wk, cp, op, 0, self.bps, self.grid_worker)
        K.evaluate(self.wk, self.sv, cp, self.b_progress.ptr + 48 * self.rank, None, 0, 0,
                   self.eval_scratch.data_ptr(), self.grid_eval)

    # ----------------------------------------------------------------------------------------
    # one pass
    # ----------------------------------------------------------------------------------------
    def use_split_protocol(self) -> None:
        """Run this engine with the split kernels (landing slots + push / version flags) instead
        of the persistent multi-rank pass.  Only before the first pass: the two keep the
        workers' copies of the KKT-filter marks differently."""
        if self.replicated:
            if self._passes_run:
                raise RuntimeError("cannot switch protocols after a pass has run")
            self.replicated = False

    def _enqueue(self, n_steps: int) -> None:
        if self.cfg.collectives == "nccl":
            from .darlin_nccl import enqueue_pass_nccl
            return enqueue_pass_nccl(self, n_steps)
        if self.cfg.fused and self.grid_fused > 0 and (self.world == 1 or self.replicated):
            K.fused_pass(self.wk, self.sv, self.ctrl_dev.data_ptr(), self.order_dev.data_ptr(),
                         self.blk_tab.data_ptr(), n_steps, self.bps, self.grid_fused,
                         self.bar.data_ptr(),
                         self.arena.peer_ptr(self.rank, self.b_progress, 6 * self.rank),
                         self.prog_peer.data_ptr(), self.world, self.eval_scratch.data_ptr(),
                         self.prof.data_ptr() if self.cfg.profile_phases else None)
        else:
            self.use_split_protocol()
            self._enqueue_split(n_steps)
        if self.world > 1:
            from .._lib import check, cuda_lib, current_stream
            check(cuda_lib().psb_rank_barrier_dev(self.bar_peers, self.rank, self.world,
                                                  self.epoch_dev.data_ptr(), current_stream()))

    def _enqueue_split(self, n_steps: int) -> None:
        K.enqueue_pass(self.wk, self.sv, self.ctrl_dev.data_ptr(), self.order_dev.data_ptr(),
                       n_steps, self.bps, self.grid_worker, self.grid_server,
                       self.arena.peer_ptr(self.rank, self.b_progress, 6 * self.rank),
                       self.prog_peer.data_ptr(), self.world, self.eval_scratch.data_ptr(),
                       self.grid_eval)

    def plan_steps(self, order: list, n_prior: int = 0) -> list:
        """Group the block order into steps of <= tau+1 blocks; prior blocks run alone
        (zero delay is forced for them, ref darlin.h:85-88)."""
        steps = [[b] for b in order[:n_prior]]
        rest = order[n_prior:]
        steps += [rest[i:i + self.bps] for i in range(0, len(rest), self.bps)]
        return steps

    def stage_pass(self, order: list, kkt_threshold: float, reset_kkt_filter: bool,
                   n_prior: int = 0) -> int:
        """Fill the pinned staging buffers and copy order + ctrl to the device (H2D)."""
        cfg = self.cfg
        steps = self.plan_steps(order, n_prior)
        n_steps = len(steps)
        flat = np.full(n_steps * self.bps, -1, dtype=np.int32)
        for i, st in enumerate(steps):
            flat[i * self.bps:i * self.bps + len(st)] = st
        self.order_host[:flat.size] = torch.from_numpy(flat)
        ctrl = K.DarlinCtrl(kkt_threshold, cfg.lambda_l1, cfg.eta, cfg.delta_max_value, self._seq,
                            int(reset_kkt_filter), n_steps, int(cfg.revive_on_reset),
                            0, int(cfg.kernel_flags), 0)
        self.ctrl_host.copy_(torch.frombuffer(bytearray(bytes(ctrl)), dtype=torch.uint8))
        self.order_dev[:flat.size].copy_(self.order_host[:flat.size], non_blocking=True)
        self.ctrl_dev.copy_(self.ctrl_host, non_blocking=True)
        self.epoch_dev.fill_(self._seq + n_steps)
        self.h2d_bytes_per_pass = flat.size * 4 + C.sizeof(K.DarlinCtrl)
        return n_steps

    def finish_pass(self, n_steps: int) -> Progress:
        """Merge the progress records of all ranks (already copied to pinned host memory)."""
        self._seq += n_steps
        self._pass_index += 1
        self._passes_run += 1
        fused = self.cfg.fused and self.grid_fused > 0 and (self.world == 1 or self.replicated)
        self.launches_per_pass = (1 if fused else 2 + 3 * n_steps) + (1 if self.world > 1 else 0)
        self.d2h_bytes_per_pass = self.prog_host.numel() * 8
        p = Progress()
        raw = self.prog_host.numpy()
        for r in range(self.world):
            rec = raw[6 * r:6 * r + 6]
            loss, pen, vio = rec[:3].view(np.float64)
            p.objective += float(loss) + float(pen)
            p.violation = max(p.violation, float(vio))
            p.nnz_w += int(rec[3])
            p.nnz_active_set += int(rec[4])
        return p

    def run_pass(self, order: list, kkt_threshold: float, reset_kkt_filter: bool,
                 n_prior: int = 0) -> Progress:
        """H2D(order, ctrl) -> device pass (graph) -> D2H(progress).  Returns merged progress."""
        cfg = self.cfg
        stream = torch.cuda.current_stream(self.device)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record(stream)
        n_steps = self.stage_pass(order, kkt_threshold, reset_kkt_filter, n_prior)
        ev[1].record(stream)
        if cfg.use_graph and cfg.collectives != "nccl":
            g = self._graphs.get(n_steps)
            if g is None:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self._side_stream()):
                    self._enqueue(n_steps)
                self._graphs[n_steps] = g
                ev[1].record(stream)
            g.replay()
        else:
            self._enqueue(n_steps)
        ev[2].record(stream)
        self.prog_host.copy_(self.b_progress.tensor, non_blocking=True)
        ev[3].record(stream)
        ev[3].synchronize()
        p = self.finish_pass(n_steps)
        p.device_ms = ev[0].elapsed_time(ev[3])
        p.kernel_ms = ev[1].elapsed_time(ev[2])
        return p

    def capture(self, n_steps: int | None = None) -> None:
        """Capture the pass graph ahead of the first pass (set-up, changes no state)."""
        if not self.cfg.use_graph or self.cfg.collectives == "nccl":
            return
        n_steps = n_steps if n_steps is not None else len(self.plan_steps(list(self.blk_order)))
        if n_steps in self._graphs:
            return
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self._side_stream()):
            self._enqueue(n_steps)
        self._graphs[n_steps] = g

    def reset_model(self) -> None:
        """Back to the zero model (init_w = ZERO): weights, margins, trust regions, active sets."""
        torch.cuda.synchronize(self.device)
        self.comm.barrier()
        self.b_w.tensor.zero_(); self.b_dw.tensor.zero_()
        self.b_land_g.tensor.zero_(); self.b_land_u.tensor.zero_()
        self.z.zero_(); self.w_cache.zero_(); self.violation.zero_(); self.acc.zero_()
        self.srv_delta.fill_(self.cfg.delta_init_value)
        self.wk_delta.fill_(self.cfg.delta_init_value)
        self.srv_active.copy_(_full_bitmap(self.n_keys, self.device))
        self.wk_active.copy_(_full_bitmap(self.n_cols, self.device))
        self.progress = []
        self._order_rng = np.random.RandomState(self.cfg.seed + 23)
        torch.cuda.synchronize(self.device)
        self.comm.barrier()

    def gather_weights(self) -> torch.Tensor:
        """The whole model in local-column order (= concatenated group dictionaries), assembled
        from every rank's shard; KKT-filtered (NaN-marked) weights read as 0."""
        parts = self.comm.gather_var(self.shard_weights().clone())
        w = torch.zeros(max(self.n_cols, 1), dtype=torch.float64, device=self.device)
        for s, ws in enumerate(parts):
            ws = ws.to(self.device)
            for gid in self.grp_ids:
                lo, hi = self.slices[gid][s]
                base, cb = self.shard_base[(gid, s)], self.col_base[gid]
                w[cb + lo:cb + hi] = ws[base:base + hi - lo]
        return torch.nan_to_num(w, nan=0.0)

    def verify(self, reported_objective: float | None = None, tol: float = 1e-9) -> dict:
        """Invariant check with plain fp64 torch (outside any timed region): the margins the
        kernels maintain incrementally must equal y_i <x_i, w> recomputed from the gathered
        shards, and the objective recomputed from them must equal the reported one."""
        w = self.gather_weights()
        counts = self.col_ptr[1:] - self.col_ptr[:-1]
        col = torch.repeat_interleave(torch.arange(self.n_cols, device=self.device), counts)
        rows = self.row_idx.long()
        sv = self.sval.double() if self.sval is not None else self.y.double()[rows]
        z_ref = torch.zeros_like(self.z).index_add_(0, rows, sv * w[col])
        del col, sv
        err = ((z_ref - self.z).abs() / (1 + z_ref.abs())).max().item() if self.n_rows else 0.0
        loss = torch.where(z_ref > 0, torch.log1p(torch.exp(-z_ref)),
                           -z_ref + torch.log1p(torch.exp(z_ref))).sum().item()
        pen = self.cfg.lambda_l1 * self.shard_weights().nan_to_num(nan=0.0).abs().sum().item()
        obj = self.comm.allreduce_sum([loss + pen], self.device)[0]
        errs = self.comm.gather_obj(err)
        out = {"max_rel_margin_error": max(errs), "objective_recomputed": obj}
        ok = max(errs) <= tol
        if reported_objective is not None:
            out["objective_rel_error"] = abs(obj - reported_objective) / max(abs(obj), 1e-300)
            ok = ok and out["objective_rel_error"] <= tol
        out["verified"] = bool(ok)
        return out

    def _side_stream(self):
        if not hasattr(self, "_cap_stream"):
            self._cap_stream = torch.cuda.Stream(self.device)
        return self._cap_stream

    # ----------------------------------------------------------------------------------------
    # DarlinScheduler::run
    # ----------------------------------------------------------------------------------------
    def next_order(self, it: int):
        order = list(self.blk_order)
        if self.cfg.random_feature_block_order:
            self._order_rng.shuffle(order)
        n_prior = 0
        if it == 0 and self.prior_blk_order:
            order = list(self.prior_blk_order) + order
            n_prior = len(self.prior_blk_order)
        return order, n_prior

    def train(self, max_pass: int | None = None, verbose: bool = True) -> list:
        cfg = self.cfg
        kkt_threshold, reset = 1e20, False
        max_iter = cfg.max_pass_of_data if max_pass is None else max_pass
        t_start = time.time()
        if verbose and self.rank == 0:
            print(f"Train l_1 logistic regression by block coordinate descent: "
                  f"{len(self.blocks)} blocks, tau={cfg.max_block_delay}", file=self.out)
        for it in range(max_iter):
            order, n_prior = self.next_order(it)
            p = self.run_pass(order, kkt_threshold, reset, n_prior)
            prev = self.progress[-1].objective if self.progress else None
            p.relative_obj = 1.0 if prev is None else prev / p.objective - 1
            p.total_time = time.time() - t_start
            p.busy_time = p.device_ms / 1e3
            self.progress.append(p)
            if verbose and self.rank == 0:
                self.show_progress(len(self.progress) - 1, kkt_threshold)
            kkt_threshold = p.violation / max(self.num_ex_global, 1) * cfg.kkt_filter_threshold_ratio
            rel = p.relative_obj
            if 0 < rel <= cfg.epsilon:
                if reset:
                    if verbose and self.rank == 0:
                        print(f"Stopped: relative objective <= {cfg.epsilon}", file=self.out)
                    break
                reset = True
            else:
                reset = False
            if it == max_iter - 1 and verbose and self.rank == 0:
                print(f"Reached maximal {max_iter} data passes", file=self.out)
        return self.progress

    def show_progress(self, it: int, kkt_threshold: float) -> None:
        """Same table as bcd.h:146-178 + darlin.h:136-156."""
        o = self.out
        if it == 0:
            print("     |        training        |  sparsity |      KKT filter     |    time (sec.)", file=o)
            print("iter |  objective    relative |     |w|_0 | threshold  #activet |(app:min max) total", file=o)
            print(" ----+------------------------+-----------+---------------------+-----------------", file=o)
        p = self.progress[it]
        ttl = p.total_time - (self.progress[it - 1].total_time if it > 0 else 0)
        print("%4d | %.5e  %.3e |%10d | %.1e %11d |%6.1f%6.1f%6.1f" % (
            it, p.objective, p.relative_obj, p.nnz_w, kkt_threshold, p.nnz_active_set,
            p.busy_time, p.busy_time, ttl), file=o)

    # ----------------------------------------------------------------------------------------
    # model access / save  (ref bcd.h:251-272)
    # ----------------------------------------------------------------------------------------
    def shard_keys(self) -> torch.Tensor:
        parts = []
        for gid in self.grp_ids:
            lo, hi = self.slices[gid][self.rank]
            parts.append(self.dicts[gid][lo:hi])
        return torch.cat(parts) if parts else torch.empty(0, dtype=torch.int64, device=self.device)

    def shard_weights(self) -> torch.Tensor:
        return self.b_w.tensor[:self.n_keys]

    def save_model(self, path_prefix: str, node_id: str | None = None) -> str:
        """Text dump `key\\tvalue` of the non-zero, non-NaN weights of this shard."""
        from ..utils.model_io import write_model_text
        node_id = node_id if node_id is not None else f"S{self.rank}"
        keys = self.shard_keys().cpu().numpy().view(np.uint64)
        w = self.shard_weights().cpu().numpy()
        return write_model_text(f"{path_prefix}_{node_id}", keys, w)

    def close(self) -> None:
        self._graphs.clear()
        if hasattr(self, "arena"):
            torch.cuda.synchronize(self.device)
            self.arena.close()


def run_pass_lockstep(engines: list, order: list, kkt_threshold: float, reset: bool,
                      n_prior: int = 0) -> Progress:
    """Drive several emulated ranks that live in ONE process on ONE GPU, phase by phase.

    Every kernel is launched only after the kernels it waits on have been enqueued on the
    same stream, so no launch ever spins on a flag that a later launch would raise (the
    rule for wait-on-each-other kernels on a single GPU).  Used by tests and debugging.
    """
    n_steps = 0
    for e in engines:
        e.use_split_protocol()
        n_steps = e.stage_pass(order, kkt_threshold, reset, n_prior)
    for e in engines:
        K.pass_begin(e.wk, e.sv, e.ctrl_dev.data_ptr())
    for i in range(n_steps):
        for e in engines:
            K.grad_push(e.wk, e.ctrl_dev.data_ptr(), e.order_dev.data_ptr(), i, e.bps, e.grid_worker)
        for e in engines:
            K.server_update(e.sv, e.ctrl_dev.data_ptr(), e.order_dev.data_ptr(), i, e.bps, e.grid_server)
        for e in engines:
            K.pull_update(e.wk, e.ctrl_dev.data_ptr(), e.order_dev.data_ptr(), i, e.bps, e.grid_worker)
    for e in engines:
        K.evaluate(e.wk, e.sv, e.ctrl_dev.data_ptr(),
                   e.arena.peer_ptr(e.rank, e.b_progress, 6 * e.rank), e.prog_peer.data_ptr(),
                   e.world, 0, e.eval_scratch.data_ptr(), e.grid_eval)
    torch.cuda.synchronize()
    out = None
    for e in engines:
        e.prog_host.copy_(e.b_progress.tensor)
        p = e.finish_pass(n_steps)
        out = out or p
    return out


def run_pass_emulated(engines: list, order: list, kkt_threshold: float, reset: bool,
                      n_prior: int = 0, grid_per_rank: int = 0) -> Progress:
    """All ranks of the persistent multi-rank pass in ONE cooperative launch on one GPU
    (psb_darlin_fused_pass_emulated): the engines live in one process (ThreadComm), their
    "peers" are pointers into the same device.  Exercises exactly the multi-GPU code path."""
    e0 = engines[0]
    R = len(engines)
    assert all(e.replicated for e in engines), "engines were switched to the split protocol"
    n_steps = 0
    for e in engines:
        n_steps = e.stage_pass(order, kkt_threshold, reset, n_prior)
    if grid_per_rank <= 0:
        grid_per_rank = max(1, min(e0.grid_fused // R, 64))
    recs = (K.DarlinRankArgs * R)()
    for r, e in enumerate(engines):
        recs[r].wk, recs[r].sv = e.wk, e.sv
        recs[r].ctrl, recs[r].order = e.ctrl_dev.data_ptr(), e.order_dev.data_ptr()
        recs[r].blk_tab, recs[r].bar = e.blk_tab.data_ptr(), e.bar.data_ptr()
        recs[r].out_local = e.arena.peer_ptr(e.rank, e.b_progress, 6 * e.rank)
        recs[r].out_peer, recs[r].scratch = e.prog_peer.data_ptr(), e.eval_scratch.data_ptr()
    raw = torch.frombuffer(bytearray(bytes(recs)), dtype=torch.uint8).to(e0.device)
    K.fused_pass_emulated(raw.data_ptr(), R, e0.sval is None, n_steps, e0.bps, grid_per_rank, R)
    torch.cuda.synchronize()
    out = None
    for e in engines:
        e.prog_host.copy_(e.b_progress.tensor)
        p = e.finish_pass(n_steps)
        out = out or p
    return out


_M64 = 0xFFFFFFFFFFFFFFFF


def _as_i64(u: int) -> int:
    u &= 0xFFFFFFFFFFFFFFFF
    return u - (1 << 64) if u >= (1 << 63) else u


def _full_bitmap(n: int, dev) -> torch.Tensor:
    words = max((n + 31) // 32, 1)
    bm = torch.full((words,), -1, dtype=torch.int32, device=dev)
    rem = n - (words - 1) * 32
    if n == 0:
        bm.zero_()
    elif rem < 32:
        bm[-1] = (1 << rem) - 1
    return bm
