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
        u64p, intp, vpp = ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(c_int), ctypes.POINTER(c_void_p)
        for name, args in (("dfb_mc_supported", [intp]), ("dfb_mc_round_size", [c_size_t, c_int, ctypes.POINTER(c_size_t)]),
                           ("dfb_vmm_create", [c_size_t, u64p, intp, vpp]), ("dfb_vmm_import", [c_int, c_size_t, u64p, vpp]),
                           ("dfb_vmm_release", [ctypes.c_ulonglong, c_void_p, c_size_t]),
                           ("dfb_mc_create", [c_int, c_size_t, u64p, intp]), ("dfb_mc_import", [c_int, u64p]),
                           ("dfb_mc_add_device", [ctypes.c_ulonglong]),
                           ("dfb_mc_bind_map", [ctypes.c_ulonglong, ctypes.c_ulonglong, c_size_t, vpp]),
                           ("dfb_mc_release", [ctypes.c_ulonglong, c_void_p, c_size_t])):
            fn = getattr(L, name)
            fn.restype, fn.argtypes = c_int, args
        L._dfb_symm_sigs = True
    return L


def multicast_supported(device: Optional[torch.device] = None) -> bool:
    """True when the device can join an NVSwitch multicast object (NVLS) and export POSIX-fd memory handles."""
    if not torch.cuda.is_available():
        return False
    out = c_int(0)
    with torch.cuda.device(device or torch.cuda.current_device()):
        rc = _sigs().dfb_mc_supported(ctypes.byref(out))
    return rc == 0 and out.value == 1


def exchange_fds(rank: int, world: int, my_fds: dict, token: str, timeout: float = 60.0) -> dict:
    """All-to-all hand-over of file descriptors between the rank processes of one box.

    `my_fds[dst]` is the list of descriptors this rank sends to `dst` (a descriptor may be sent to several peers).
    Returns `{src: [fds...]}` with the descriptors as they are numbered in THIS process.  Transport: abstract-namespace
    unix sockets (a leading NUL byte, then `dfb-<token>-<rank>`, nothing touches the filesystem) and SCM_RIGHTS; every rank listens first
    (callers put a barrier between `listen` and `send` by way of the token broadcast), then sends, then collects.
    """
    import socket
    import struct
    import time
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(f"\0dfb-{token}-{rank}")
    srv.listen(world + 1)
    srv.settimeout(timeout)
    try:
        for dst in range(world):
            if dst == rank:
                continue
            fds = list(my_fds.get(dst, []))
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            deadline = time.time() + timeout
            while True:                      # the peer may not be listening yet
                try:
                    c.connect(f"\0dfb-{token}-{dst}")
                    break
                except (ConnectionRefusedError, FileNotFoundError):
                    if time.time() > deadline:
                        raise TimeoutError(f"rank {rank}: peer {dst} never opened its descriptor socket")
                    time.sleep(0.01)
            socket.send_fds(c, [struct.pack("<ii", rank, len(fds))], fds)
            c.close()
        got: dict = {}
        for _ in range(world - 1):
            conn, _addr = srv.accept()
            conn.settimeout(timeout)
            msg, fds, _flags, _a = socket.recv_fds(conn, 8, 16)
            src, n = struct.unpack("<ii", msg)
            if n != len(fds):
                raise RuntimeError(f"rank {rank}: expected {n} descriptors from rank {src}, got {len(fds)}")
            got[src] = list(fds)
            conn.close()
        return got
    finally:
        srv.close()


class _RawCuda:
    """Exposes a raw device pointer through `__cuda_array_interface__` so torch can view it without a copy."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2,
                                         "strides": None}
        self._owner = owner


class SymmetricHeap:
    def __init__(self, nbytes: int, rank: int = 0, world: int = 1, group=None, device: Optional[torch.device] = None,
                 _emulated_peers: Optional[List["SymmetricHeap"]] = None, multicast: bool = False):
        L = _sigs()
        if world > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks per box")
        self.rank, self.world, self.group = rank, world, group
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.nbytes = int((nbytes + PAD_BYTES + 4095) // 4096 * 4096)
        self.mc_base = 0                   # multicast (NVLS) address of the heap, 0 when the heap is unicast only
        self._vmm = None
        if multicast and world > 1 and _emulated_peers is None:
            self._init_multicast()
            return
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

    def _init_multicast(self) -> None:
        """VMM allocation + peer unicast mappings + one NVSwitch multicast mapping (csrc/mcast.cu)."""
        import os
        import torch.distributed as dist
        L = _sigs()
        u64 = ctypes.c_ulonglong
        with torch.cuda.device(self.device):
            rounded = c_size_t(0)
            check(L.dfb_mc_round_size(self.nbytes, self.world, ctypes.byref(rounded)), "dfb_mc_round_size")
            self.nbytes = int(rounded.value)
            h, fd, va = u64(0), c_int(-1), c_void_p()
            check(L.dfb_vmm_create(self.nbytes, ctypes.byref(h), ctypes.byref(fd), ctypes.byref(va)), "dfb_vmm_create")
            mc, mc_fd = u64(0), c_int(-1)
            if self.rank == 0:
                check(L.dfb_mc_create(self.world, self.nbytes, ctypes.byref(mc), ctypes.byref(mc_fd)), "dfb_mc_create")
            tok = [f"{os.getpid()}-{int.from_bytes(os.urandom(4), 'little')}" if self.rank == 0 else None]
            dist.broadcast_object_list(tok, src=0, group=self.group)
            send = {d: ([fd.value] + ([mc_fd.value] if self.rank == 0 else [])) for d in range(self.world) if d != self.rank}
            got = exchange_fds(self.rank, self.world, send, tok[0])
            self.base = int(va.value)
            self.peer_bases = [0] * self.world
            self.peer_bases[self.rank] = self.base
            imported = []
            for src, fds in got.items():
                ph, pva = u64(0), c_void_p()
                check(L.dfb_vmm_import(fds[0], self.nbytes, ctypes.byref(ph), ctypes.byref(pva)), f"dfb_vmm_import(rank {src})")
                self.peer_bases[src] = int(pva.value)
                imported.append((ph.value, int(pva.value)))
                if src == 0:
                    check(L.dfb_mc_import(fds[1], ctypes.byref(mc)), "dfb_mc_import")
                for f in fds:
                    os.close(f)
            os.close(fd.value)
            if self.rank == 0:
                os.close(mc_fd.value)
            check(L.dfb_mc_add_device(mc), "dfb_mc_add_device")
            dist.barrier(group=self.group)                      # every device is in the group before anybody binds
            mva = c_void_p()
            check(L.dfb_mc_bind_map(mc, h, self.nbytes, ctypes.byref(mva)), "dfb_mc_bind_map")
            dist.barrier(group=self.group)
        self.mc_base = int(mva.value)
        self._vmm = {"handle": h.value, "mc": mc.value, "imported": imported}
        self._cursor = PAD_BYTES
        self._bytes = torch.as_tensor(_RawCuda(self.base, self.nbytes, self), device=self.device)
        self._opened = []
        self._emulated = False

    def mc_ptr(self, offset: int = 0) -> c_void_p:
        """Multicast address of `offset` (0 when the heap has no NVLS mapping)."""
        return c_void_p(self.mc_base + offset if self.mc_base else 0)

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

    # ---- failure detection ----------------------------------------------------------------------
    def set_barrier_timeout(self, seconds: float) -> None:
        """Bound of every in-kernel wait for a peer on this heap's device (abstract_client.ts:12-13: the reference's
        connect / upload time-outs).  A kernel whose peer stays silent that long records the peer and finishes."""
        from ..ops._lib import lib
        L = lib()
        L.dfb_set_barrier_timeout_us.restype = c_int
        L.dfb_set_barrier_timeout_us.argtypes = [ctypes.c_int64, c_void_p]
        with torch.cuda.device(self.device):
            check(L.dfb_set_barrier_timeout_us(int(seconds * 1e6), c_void_p(torch.cuda.current_stream(self.device).cuda_stream)),
                  "dfb_set_barrier_timeout_us")

    def barrier_failure(self, stream: Optional["torch.cuda.Stream"] = None) -> Optional[tuple]:
        """(silent rank, barrier generation) if a kernel of this rank gave up waiting since the last call, else None.
        `stream`: pass a side stream to ask while the main stream is still busy (the watchdog thread does)."""
        from ..ops._lib import lib
        L = lib()
        L.dfb_barrier_failures.restype = c_int
        L.dfb_barrier_failures.argtypes = [c_void_p, c_void_p, c_void_p]
        out = (ctypes.c_uint * 2)()
        st = stream or torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            check(L.dfb_barrier_failures(self.pad_ptr(), out, c_void_p(st.cuda_stream)), "dfb_barrier_failures")
        return (int(out[0]) - 1, int(out[1])) if out[0] else None

    def pad_ptr(self, rank: Optional[int] = None) -> c_void_p:
        return c_void_p(self.peer_bases[self.rank if rank is None else rank])

    def close(self) -> None:
        L = _sigs()
        if self._vmm is not None:
            v, self._vmm = self._vmm, None
            self._bytes = None
            L.dfb_mc_release(v["mc"], c_void_p(self.mc_base), self.nbytes)
            for ph, pva in v["imported"]:
                L.dfb_vmm_release(ph, c_void_p(pva), self.nbytes)
            L.dfb_vmm_release(v["handle"], c_void_p(self.base), self.nbytes)
            self.base = self.mc_base = 0
            return
        for p in self._opened:
            L.dfb_symm_close_handle(c_void_p(p))
        self._opened = []
        if self.base:
            self._bytes = None
            L.dfb_symm_free(c_void_p(self.base))
            self.base = 0
