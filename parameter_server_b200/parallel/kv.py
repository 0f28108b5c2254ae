"""DeviceKVVector: the key-range-partitioned KV container of the public API, on the device.

ref: KVVector / SharedParameter (src/parameter/kv_vector.h, shared_parameter.h).  Every rank owns
the shard of a global sorted key dictionary that falls into its key range (even split of the
observed key range, or Range::all()).  `register(keys)` resolves a sorted key list to (owner, slot)
once — the device form of sliceKeyOrderedMsg + the KEY_CACHING filter — and returns a handle;
`push(handle, values)` adds into the owners' arrays and `pull(handle)` gathers from them, both as
single kernels that touch peer HBM directly (NVLink), with no NCCL and no host staging.

Consistency: pushes and pulls are ordered on the calling stream; `fence()` is a cross-rank device
barrier (all earlier pushes of all ranks are visible afterwards) — the bulk-synchronous subset of
the reference's timestamp / wait_time protocol, which the learners implement with per-step flags.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from .._lib import check, cuda_lib, current_stream, ptr
from ..ops import localizer as L
from ..ops._sig import apply_cuda_sigs
from ..utils.range import Range
from .comm import Comm, SingleComm


@dataclass
class KVHandle:
    keys: torch.Tensor
    owner: torch.Tensor
    slot: torch.Tensor
    n: int
    signature: int
    # bulk path (register(..., bulk=True)): owner segments of my keys and, as an owner, the (32-bit)
    # slot lists of every worker's segment for me and where that segment starts in its staging
    seg_begin: list | None = None
    slots_from: list | None = None
    src_off_from: list | None = None


class DeviceKVVector:
    def __init__(self, dictionary: torch.Tensor, comm: Comm | None = None, device=None,
                 shard_by: str = "data_range", bulk_capacity: int = 0):
        """`dictionary`: the global sorted unique key list (identical on every rank)."""
        apply_cuda_sigs()
        self.lib = cuda_lib()
        self.comm = comm or SingleComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.dict = dictionary.to(self.device).contiguous()
        n = self.dict.numel()
        W = self.world
        if shard_by == "all" or n == 0:
            whole = Range.all()
        else:
            u = lambda t: int(t.item()) & 0xFFFFFFFFFFFFFFFF  # noqa: E731
            whole = Range(u(self.dict[0]), u(self.dict[-1]) + 1)
        ranges = [whole.even_divide(W, s) for s in range(W)]
        probes = torch.tensor([_as_i64(r.begin) for r in ranges[1:]], dtype=torch.int64, device=self.device)
        cut = L.lower_bound(self.dict, probes).tolist() if W > 1 and n else [n] * (W - 1)
        begins = [0] + cut + [n]
        self.shard_begin = torch.tensor(begins, dtype=torch.int64, device=self.device)
        self.shard_len = [begins[s + 1] - begins[s] for s in range(W)]
        cap = (max(max(self.shard_len), 1) + 3) // 4 * 4
        # bulk path: this rank's push staging (values sorted by owner, read by the owners over
        # NVLink) and per (worker, owner) landing regions for pulls
        self.bulk_cap = (int(bulk_capacity) + 3) // 4 * 4
        self.arena = self.comm.make_arena(cap * 4 + 2 * W * self.bulk_cap * 4 + 3 * W * 8 + 8192, self.device)
        self.b_val = self.arena.alloc("val", cap, torch.float32)
        self.b_bar = self.arena.alloc("bar", W, torch.int64)
        self.b_land = self.arena.alloc("push_staging", W * self.bulk_cap, torch.float32)
        self.b_pull = self.arena.alloc("land_pull", W * self.bulk_cap, torch.float32)
        self.b_pflag = self.arena.alloc("push_flag", W, torch.int64)
        self.b_gflag = self.arena.alloc("pull_flag", W, torch.int64)
        self.ctr = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.apply_bar = torch.zeros(4, dtype=torch.int32, device=self.device)
        self._seq = 0
        self.peers = (C.c_void_p * W)(*[self.arena.peer_ptr(s, self.b_val) for s in range(W)])
        self.bar_peers = (C.c_void_p * W)(*[self.arena.peer_ptr(s, self.b_bar) for s in range(W)])
        self._epoch = 0
        torch.cuda.synchronize(self.device)
        self.comm.barrier()

    @property
    def value(self) -> torch.Tensor:
        """This rank's shard values (aligned with my slice of the dictionary)."""
        return self.b_val.tensor[:self.shard_len[self.rank]]

    @property
    def key(self) -> torch.Tensor:
        a = int(self.shard_begin[self.rank])
        return self.dict[a:a + self.shard_len[self.rank]]

    def register(self, keys: torch.Tensor, bulk: bool = False) -> KVHandle:
        keys = keys.to(self.device).contiguous()
        n = keys.numel()
        idx1 = L.lookup_sorted(self.dict, keys)
        owner = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)
        slot = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        check(self.lib.psb_kv_route_dense(ptr(idx1), n, ptr(self.shard_begin), self.world, ptr(owner),
                                          ptr(slot), current_stream()))
        assert n == 0 or int(slot[:n].min().item()) >= 0, "keys outside the dictionary"
        h = KVHandle(keys, owner, slot, n, L.key_signature(keys) if n else 0)
        assert max(self.shard_len) < 2 ** 31, "bulk slots are 32 bit"
        if bulk:  # collective: every rank registers its key list at the same time
            W = self.world
            counts = torch.bincount(owner[:n].long(), minlength=W).tolist() if n else [0] * W
            seg = [0]
            for c in counts:
                seg.append(seg[-1] + c)
            assert max(counts) <= self.bulk_cap, "bulk_capacity too small for this key list"
            h.seg_begin = seg
            # the slot lists travel once (key caching); afterwards only values move
            h.slots_from = [t.to(torch.int32).contiguous() for t in
                            self.comm.alltoall_var([slot[seg[s]:seg[s + 1]].contiguous() for s in range(W)])]
            h.src_off_from = [sb[self.rank] for sb in self.comm.gather_obj(seg)]
        return h

    def push(self, h: KVHandle, values: torch.Tensor) -> None:
        assert values.dtype == torch.float32 and values.numel() == h.n
        check(self.lib.psb_kv_push_add_f32(ptr(values), ptr(h.owner), ptr(h.slot), h.n, self.peers,
                                           self.world, current_stream()))

    def pull(self, h: KVHandle, out: torch.Tensor | None = None) -> torch.Tensor:
        if out is None:
            out = torch.empty(max(h.n, 4), dtype=torch.float32, device=self.device)
        check(self.lib.psb_kv_pull_f32(ptr(out), ptr(h.owner), ptr(h.slot), h.n, self.peers, self.world,
                                       current_stream()))
        return out[:h.n]

    # ---- bulk path (collective calls: every rank pushes / pulls its handle) ----------------------
    def push_buffer(self, h: KVHandle) -> torch.Tensor:
        """The staging view to fill for a zero-copy push_bulk (values in key order); valid until the
        next collective call."""
        return self.b_land.tensor[:h.n]

    def push_bulk(self, h: KVHandle, values: torch.Tensor) -> None:
        """Receiver-driven: my values stay in my staging buffer; after a rank barrier every owner
        adds each worker's segment into its shard, reading it over NVLink inside the apply kernel
        (one cooperative launch, fixed worker order: deterministic).  Collective."""
        self.fence()  # nobody still reads the staging of the previous collective call
        self._bulk_stage(h, values)
        self.fence()  # every rank's staging is complete
        self._push_apply(h)

    def _bulk_stage(self, h: KVHandle, values: torch.Tensor) -> None:
        assert values.dtype == torch.float32 and values.numel() == h.n
        st = self.push_buffer(h)
        if values.data_ptr() != st.data_ptr():
            st.copy_(values)

    def _push_apply(self, h: KVHandle) -> None:
        W, A = self.world, self.arena
        src = (C.c_void_p * W)(*[A.peer_ptr(w, self.b_land, h.src_off_from[w]) for w in range(W)])
        slots = (C.c_void_p * W)(*[ptr(h.slots_from[w]) for w in range(W)])
        counts = (C.c_int64 * W)(*[h.slots_from[w].numel() for w in range(W)])
        check(self.lib.psb_kv_bulk_pull_apply(self.b_val.ptr, src, slots, counts, W, self.rank,
                                              ptr(self.apply_bar), current_stream()))

    def pull_bulk(self, h: KVHandle) -> torch.Tensor:
        """Owner side gathers every worker's slots and stores them contiguously into that worker's
        buffer; the result (my keys' values, in key order) is a view of my local landing buffer.
        Collective."""
        self.fence()
        self._bulk_gather_send(h)
        return self._bulk_pull_result(h)

    def _bulk_gather_send(self, h: KVHandle) -> None:
        W, A, r = self.world, self.arena, self.rank
        self._seq += 1
        slots = (C.c_void_p * W)(*[ptr(h.slots_from[w]) for w in range(W)])
        counts = (C.c_int64 * W)(*[h.slots_from[w].numel() for w in range(W)])
        dst = (C.c_void_p * W)(*[A.peer_ptr(w, self.b_pull, r * self.bulk_cap) for w in range(W)])
        flg = (C.c_void_p * W)(*[A.peer_ptr(w, self.b_gflag, r) for w in range(W)])
        check(self.lib.psb_kv_bulk_gather_send32(self.b_val.ptr, slots, counts, dst, flg, W, self.rank,
                                                 self._seq, ptr(self.ctr), current_stream()))

    def _bulk_pull_result(self, h: KVHandle) -> torch.Tensor:
        W = self.world
        check(self.lib.psb_kv_wait_flags(self.b_gflag.ptr, W, self._seq, current_stream()))
        parts = [self.b_pull.tensor[s * self.bulk_cap:s * self.bulk_cap + h.seg_begin[s + 1] - h.seg_begin[s]]
                 for s in range(W)]
        return parts[0] if W == 1 else torch.cat(parts)

    def fence(self) -> None:
        """Cross-rank barrier on the device (flags over peer memory)."""
        if self.world == 1:
            return
        self._epoch += 1
        check(self.lib.psb_rank_barrier(self.bar_peers, self.rank, self.world, self._epoch, current_stream()))

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        self.arena.close()


def _as_i64(u: int) -> int:
    u &= 0xFFFFFFFFFFFFFFFF
    return u - (1 << 64) if u >= (1 << 63) else u
