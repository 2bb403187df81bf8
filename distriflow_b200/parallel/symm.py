"""Symmetric heap: one device allocation per rank, mapped into every peer, same offsets everywhere.

This is the B200 replacement of the reference's transport (socket.io star): ranks of
one NVSwitch box exchange `cudaIpcMemHandle`s once (bootstrap through `torch.distributed`, any backend)
and afterwards kernels load/store peer memory directly.  The first `PAD_BYTES` of every heap are the
signal pad (barrier flags, async control block, work-queue words; layout in `csrc/comm.cu`).

`SymmetricHeap.emulated(R, nbytes)` builds R heaps inside one process on one GPU so the collective
kernels can be exercised rank by rank without co-scheduled peers (B200_PROFILING.md: "emulate the ranks
... or test the multi-rank path on the CPU").
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from ..ops import _lib
from ..ops._lib import c_int, c_size_t, c_void_p, check

PAD_BYTES = 4096
MAX_RANKS = 8


def _sigs():
    L = _lib.lib()
    if not getattr(L, "_dfb_symm_sigs", False):
        L.dfb_symm_alloc.restype = c_int
        L.dfb_symm_alloc.argtypes = [c_size_t, ctypes.POINTER(c_void_p)]
        L.dfb_symm_free.restype = c_int
        L.dfb_symm_free.argtypes = [c_void_p]
        L.dfb_symm_get_handle.restype = c_int
        L.dfb_symm_get_handle.argtypes = [c_void_p, ctypes.c_char_p]
        L.dfb_symm_open_handle.restype = c_int
        L.dfb_symm_open_handle.argtypes = [ctypes.c_char_p, ctypes.POINTER(c_void_p)]
        L.dfb_symm_close_handle.restype = c_int
        L.dfb_symm_close_handle.argtypes = [c_void_p]
        L._dfb_symm_sigs = True
    return L


class _RawCuda:
    """Exposes a raw device pointer through `__cuda_array_interface__` so torch can view it without a copy."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2,
                                         "strides": None}
        self._owner = owner


class SymmetricHeap:
    def __init__(self, nbytes: int, rank: int = 0, world: int = 1, group=None, device: Optional[torch.device] = None,
                 _emulated_peers: Optional[List["SymmetricHeap"]] = None):
        L = _sigs()
        if world > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks per box")
        self.rank, self.world, self.group = rank, world, group
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.nbytes = int((nbytes + PAD_BYTES + 4095) // 4096 * 4096)
        base = c_void_p()
        with torch.cuda.device(self.device):
            check(L.dfb_symm_alloc(self.nbytes, ctypes.byref(base)), "dfb_symm_alloc")
        self.base = int(base.value)
        self._cursor = PAD_BYTES
        self._bytes = torch.as_tensor(_RawCuda(self.base, self.nbytes, self), device=self.device)
        self.peer_bases: List[int] = [self.base]
        self._opened: List[int] = []
        self._emulated = _emulated_peers is not None
        if not self._emulated and world > 1:
            self._exchange()
        elif not self._emulated:
            self.peer_bases = [self.base]

    # ---- construction -------------------------------------------------------------------------
    @classmethod
    def emulated(cls, world: int, nbytes: int, device: Optional[torch.device] = None) -> List["SymmetricHeap"]:
        heaps = [cls(nbytes, r, world, None, device, _emulated_peers=[]) for r in range(world)]
        bases = [h.base for h in heaps]
        for h in heaps:
            h.peer_bases = list(bases)
        return heaps

    def _exchange(self) -> None:
        import torch.distributed as dist
        L = _sigs()
        buf = ctypes.create_string_buffer(64)
        check(L.dfb_symm_get_handle(c_void_p(self.base), buf), "dfb_symm_get_handle")
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, bytes(buf.raw), group=self.group)
        self.peer_bases = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.peer_bases.append(self.base)
                continue
            p = c_void_p()
            with torch.cuda.device(self.device):
                check(L.dfb_symm_open_handle(ctypes.c_char_p(h), ctypes.byref(p)), f"dfb_symm_open_handle(rank {r})")
            self.peer_bases.append(int(p.value))
            self._opened.append(int(p.value))
        dist.barrier(group=self.group)

    # ---- allocation ---------------------------------------------------------------------------
    def alloc(self, nbytes: int, align: int = 256) -> int:
        """Bump allocation; every rank must issue the same sequence of calls (offsets are then identical)."""
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError(f"symmetric heap exhausted: need {off + nbytes}, have {self.nbytes}")
        self._cursor = off + nbytes
        return off

    def view(self, offset: int, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for d in shape:
            n *= int(d)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return self._bytes[offset:offset + nbytes].view(dtype).view(*shape)

    def ptrs(self, offset: int = 0):
        """ctypes array of `world` pointers: the same offset inside every rank's heap."""
        arr = (c_void_p * MAX_RANKS)()
        for r in range(self.world):
            arr[r] = self.peer_bases[r] + offset
        return arr

    @property
    def pad(self) -> torch.Tensor:
        return self.view(0, (PAD_BYTES // 4,), torch.int32)

    def pad_ptr(self, rank: Optional[int] = None) -> c_void_p:
        return c_void_p(self.peer_bases[self.rank if rank is None else rank])

    def close(self) -> None:
        L = _sigs()
        for p in self._opened:
            L.dfb_symm_close_handle(c_void_p(p))
        self._opened = []
        if self.base:
            self._bytes = None
            L.dfb_symm_free(c_void_p(self.base))
            self.base = 0
