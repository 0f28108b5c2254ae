"""Host-side collectives used OUTSIDE the hot path (preprocessing, set-up, progress merge).

Three interchangeable back ends:
  SingleComm   one rank
  TorchComm    torch.distributed (NCCL on GPU boxes, gloo in CPU tests)
  ThreadComm   several emulated ranks inside one process (threads), used to exercise the
               multi-rank index arithmetic and the peer-pointer wiring on a single GPU: the
               "peers" are then plain device pointers of the same process.

The push / pull data path never goes through these: it is peer loads/stores in the kernels.
"""
from __future__ import annotations

import threading

import torch

from .symm import SymmetricArena


class Comm:
    rank = 0
    world = 1

    def gather_var(self, t: torch.Tensor) -> list:
        return [t]

    def gather_obj(self, obj) -> list:
        return [obj]

    def allreduce_sum(self, vals: list, device=None) -> list:
        return list(vals)

    def barrier(self) -> None:
        pass

    def alltoall_var(self, parts: list) -> list:
        """parts[r] goes to rank r; returns what every rank sent to me (set-up time only)."""
        return list(parts)

    def make_arena(self, nbytes: int, device) -> SymmetricArena:
        return SymmetricArena(nbytes, 0, 1, None, device)


class SingleComm(Comm):
    pass


class TorchComm(Comm):
    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)

    def gather_var(self, t):
        if self.world == 1:
            return [t]
        n = torch.tensor([t.numel()], dtype=torch.int64, device=t.device)
        ns = [torch.zeros_like(n) for _ in range(self.world)]
        self.dist.all_gather(ns, n, group=self.group)
        sizes = [int(x.item()) for x in ns]
        pad = torch.zeros(max(max(sizes), 1), dtype=t.dtype, device=t.device)
        pad[:t.numel()] = t
        outs = [torch.empty_like(pad) for _ in range(self.world)]
        self.dist.all_gather(outs, pad, group=self.group)
        return [o[:k] for o, k in zip(outs, sizes)]

    def gather_obj(self, obj):
        out = [None] * self.world
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def allreduce_sum(self, vals, device=None):
        if self.world == 1:
            return list(vals)
        t = torch.tensor(vals, dtype=torch.float64, device=device)
        self.dist.all_reduce(t, group=self.group)
        return t.tolist()

    def barrier(self):
        self.dist.barrier(group=self.group)

    def alltoall_var(self, parts):
        dev, dtype = parts[0].device, parts[0].dtype
        send = torch.tensor([p.numel() for p in parts], dtype=torch.int64, device=dev)
        recv = torch.empty_like(send)
        self.dist.all_to_all_single(recv, send, group=self.group)
        out_split = recv.tolist()
        buf = torch.empty(sum(out_split), dtype=dtype, device=dev)
        self.dist.all_to_all_single(buf, torch.cat(parts), out_split, send.tolist(), group=self.group)
        return list(torch.split(buf, out_split))

    def make_arena(self, nbytes, device):
        return SymmetricArena(nbytes, self.rank, self.world, self.group, device)


class _ThreadShared:
    def __init__(self, world: int):
        self.world = world
        self.bar = threading.Barrier(world)
        self.slots = [None] * world


class ThreadComm(Comm):
    """Emulated ranks: create with ThreadComm.create(world) and use comm[r] in thread r."""

    def __init__(self, shared: _ThreadShared, rank: int):
        self.sh, self.rank, self.world = shared, rank, shared.world

    @staticmethod
    def create(world: int) -> list:
        sh = _ThreadShared(world)
        return [ThreadComm(sh, r) for r in range(world)]

    def gather_obj(self, obj):
        self.sh.slots[self.rank] = obj
        self.sh.bar.wait()
        out = list(self.sh.slots)
        self.sh.bar.wait()
        return out

    def gather_var(self, t):
        return self.gather_obj(t)

    def allreduce_sum(self, vals, device=None):
        rows = self.gather_obj(list(vals))
        return [sum(r[i] for r in rows) for i in range(len(vals))]

    def barrier(self):
        self.sh.bar.wait()

    def alltoall_var(self, parts):
        everyone = self.gather_obj(parts)
        return [everyone[r][self.rank] for r in range(self.world)]

    def make_arena(self, nbytes, device):
        a = SymmetricArena(nbytes, self.rank, 1, None, device)
        a.world = self.world
        bases = self.gather_obj(a.base)
        a.peer_base = list(bases)
        return a


def run_threads(world: int, fn):
    """Run fn(comm) on `world` emulated ranks; returns the list of results (re-raises errors)."""
    comms = ThreadComm.create(world)
    res, errs = [None] * world, []

    def body(r):
        try:
            res[r] = fn(comms[r])
        except BaseException as e:  # noqa: BLE001 - surfaced below
            errs.append(e)
            comms[r].sh.bar.abort()

    ts = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errs:
        raise errs[0]
    return res
