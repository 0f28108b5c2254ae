"""Symmetric peer-mapped device arena: the data plane that replaces the reference's Van.

Every rank cudaMalloc's one arena of identical size, exports it with CUDA IPC and maps all
peers' arenas (NVLink P2P through NVSwitch).  Sub-buffers are carved at identical offsets on
every rank, so `peer_ptr(r, buf) = peer_base[r] + buf.offset` — the registration step that
the reference performs per message (ZeroMQ frames, src/system/van.cc:117-223) happens once.

Handle exchange uses torch.distributed (any backend: NCCL on the GPU box, gloo in CPU tests
where only the offset arithmetic is exercised).
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass

import torch

from .._lib import check, cuda_lib
from ..ops._sig import apply_cuda_sigs

_DT_SIZE = {torch.float64: 8, torch.int64: 8, torch.float32: 4, torch.int32: 4, torch.uint8: 1,
            torch.int8: 1, torch.int16: 2, torch.float16: 2, torch.bfloat16: 2}


class _CudaArray:
    """__cuda_array_interface__ shim so torch can view raw arena memory."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


@dataclass
class SymmBuffer:
    name: str
    offset: int
    nbytes: int
    tensor: torch.Tensor  # local view

    @property
    def ptr(self) -> int:
        return self.tensor.data_ptr()


DEVICE_PAGE = 2 << 20


def arena_bytes(nbytes: int) -> int:
    """Size of the allocation behind an arena of `nbytes`: whole device pages (see SymmetricArena)."""
    granule = int(os.environ.get("PSB_ARENA_GRANULE", DEVICE_PAGE))
    return max(granule, (int(nbytes) + granule - 1) // granule * granule)


class SymmetricArena:
    def __init__(self, nbytes: int, rank: int = 0, world: int = 1, group=None, device=None):
        apply_cuda_sigs()
        self.lib = cuda_lib()
        self.rank, self.world, self.group = rank, world, group
        self.device = torch.device(device if device is not None else
                                   f"cuda:{torch.cuda.current_device()}")
        # The size MUST be a multiple of the 2 MB device page.  Measured on B200 (benchmarks/micro/
        # peer_random.cu, benchmarks/peer_table_probe.py; profiles/README.md section 0.5): a cudaMalloc of
        # 2 GB + 4 KB, exported through cudaIpc, costs a peer 1 320 us for 200 000 random 8-byte reads,
        # the same allocation rounded to 2 MB 41 us; coalesced bulk transfers do not show the difference.
        # PSB_ARENA_GRANULE exists only to reproduce that measurement.
        self.nbytes = arena_bytes(nbytes)
        base = ctypes.c_void_p(0)
        with torch.cuda.device(self.device):
            check(self.lib.psb_malloc(ctypes.byref(base), self.nbytes))
            check(self.lib.psb_memset_async(base, 0, self.nbytes, None))
            torch.cuda.synchronize()
        self.base = base.value
        self._cursor = 0
        self.buffers: dict[str, SymmBuffer] = {}
        self.peer_base = [self.base] * world
        self._mapped: list[int] = []
        self._whole = torch.as_tensor(_CudaArray(self.base, self.nbytes, self), device=self.device)
        if world > 1:
            self._exchange()

    # -- IPC -------------------------------------------------------------------------------
    def _exchange(self) -> None:
        import torch.distributed as dist
        n = self.lib.psb_ipc_handle_bytes()
        buf = (ctypes.c_uint8 * n)()
        check(self.lib.psb_ipc_get_handle(ctypes.c_void_p(self.base), buf))
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(buf), group=self.group)
        for r, h in enumerate(handles):
            if r == self.rank:
                continue
            out = ctypes.c_void_p(0)
            hb = (ctypes.c_uint8 * n).from_buffer_copy(h)
            check(self.lib.psb_ipc_open_handle(hb, ctypes.byref(out)))
            self.peer_base[r] = out.value
            self._mapped.append(out.value)
        dist.barrier(group=self.group)

    # -- carving -----------------------------------------------------------------------------
    def alloc(self, name: str, numel: int, dtype: torch.dtype, align: int = 256) -> SymmBuffer:
        """Carve `numel` elements; every rank must call with identical arguments in the same order."""
        size = _DT_SIZE[dtype] * int(numel)
        off = (self._cursor + align - 1) // align * align
        if off + size > self.nbytes:
            raise MemoryError(f"symmetric arena exhausted: need {off + size} of {self.nbytes} bytes")
        self._cursor = off + size
        t = self._whole[off:off + size].view(dtype) if size else torch.empty(0, dtype=dtype,
                                                                          device=self.device)
        b = SymmBuffer(name, off, size, t)
        self.buffers[name] = b
        return b

    def peer_ptr(self, r: int, buf: SymmBuffer, elem_offset: int = 0) -> int:
        return self.peer_base[r] + buf.offset + elem_offset * buf.tensor.element_size()

    def close(self) -> None:
        for m in self._mapped:
            self.lib.psb_ipc_close_handle(ctypes.c_void_p(m))
        self._mapped = []
        if self.base:
            self._whole = None
            for b in self.buffers.values():
                b.tensor = None
            self.lib.psb_free(ctypes.c_void_p(self.base))
            self.base = 0
