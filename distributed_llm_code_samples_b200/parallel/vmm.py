"""EXPERIMENTAL: symmetric blocks on cuMem VMM allocations with an NVLS multicast mapping (``csrc/vmm.cu``).

Same peer-table contract as ``symm.SymmBuffer`` (CUDA IPC), plus ``mc_ptr``: a multicast address on which
``multimem.ld_reduce`` / ``multimem.st`` reduce and broadcast inside the NVSwitch.  File descriptors travel between the
ranks of the node over unix datagram sockets (``SCM_RIGHTS``).  Validated on B200 in round 2 (``profiles/tests/``):
``tests/test_multigpu.py::test_vmm_multicast_all_reduce``; ``SymmWorld.alloc(..., multicast=True)`` hands these blocks to
the fused MP trainer, whose all-reduce then sums inside the switch.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native as nv


def supported() -> tuple[bool, bool]:
    """(posix-fd export, multicast) support of the current device."""
    out = (C.c_int * 2)()
    if nv.load().b200_vmm_query(out) != 0:
        return False, False
    return bool(out[0]), bool(out[1])


def _sock_path(tag: str, rank: int) -> str:
    return os.path.join(tempfile.gettempdir(), f"b200ffn_{tag}_{rank}.sock")


class VmmBuffer:
    """One shareable block per rank, every peer's block mapped locally, optional multicast mapping of all of them."""

    def __init__(self, rank: int, world: int, device: torch.device, nbytes: int, multicast: bool = True,
                 group: Optional[dist.ProcessGroup] = None, tag: str = "blk"):
        lib = nv.load()
        torch.cuda.set_device(device)
        self.rank, self.world, self.device = rank, world, device
        gran = int(lib.b200_vmm_granularity(world, int(multicast)))
        if gran <= 0:
            raise nv.NativeError("VMM allocation granularity query failed: " + lib.b200_last_error().decode())
        self.nbytes = (nbytes + gran - 1) // gran * gran
        # unique per job on this host: ranks of one job share MASTER_PORT; a world of one uses its pid
        tag = f"{tag}{os.environ.get('MASTER_PORT', '') if world > 1 else os.getpid()}"
        ptr, handle, fd = C.c_void_p(), C.c_ulonglong(), C.c_int()
        nv._check(lib.b200_vmm_alloc(self.nbytes, C.byref(ptr), C.byref(handle), C.byref(fd)), "b200_vmm_alloc")
        self.ptr, self.handle = ptr.value, handle.value
        self.peer_ptrs: List[int] = [0] * world
        self.peer_handles: List[int] = [0] * world
        self.peer_ptrs[rank] = self.ptr
        # every rank listens, then sends its fd to every peer (tag = sender rank), then receives world-1 fds
        sock = lib.b200_fd_listen(_sock_path(tag, rank).encode())
        if sock < 0:
            raise nv.NativeError(lib.b200_last_error().decode())
        if world > 1:
            dist.barrier(group=group)
        for q in range(world):
            if q != rank:
                nv._check(lib.b200_fd_send(_sock_path(tag, q).encode(), fd.value, rank), "b200_fd_send")
        early_mc_fd = None             # datagrams from different senders are not ordered: rank 0's multicast fd (tag -1)
        got = 0                        # may arrive before the last peer's block fd
        while got < world - 1:
            src = C.c_int()
            rfd = lib.b200_fd_recv(sock, C.byref(src))
            if rfd < 0:
                raise nv.NativeError(lib.b200_last_error().decode())
            if src.value == -1:
                early_mc_fd = rfd
                continue
            got += 1
            p, h = C.c_void_p(), C.c_ulonglong()
            nv._check(lib.b200_vmm_import(rfd, self.nbytes, C.byref(p), C.byref(h)), "b200_vmm_import")
            lib.b200_fd_close(rfd)
            self.peer_ptrs[src.value], self.peer_handles[src.value] = p.value, h.value
        lib.b200_fd_close(fd.value)
        self.peers_c = (C.c_void_p * world)(*self.peer_ptrs)
        self.mc_ptr: Optional[int] = None
        if multicast and world > 1:
            mc = C.c_ulonglong()
            if rank == 0:
                mfd = C.c_int()
                nv._check(lib.b200_mc_create(world, self.nbytes, C.byref(mc), C.byref(mfd)), "b200_mc_create")
                for q in range(1, world):
                    nv._check(lib.b200_fd_send(_sock_path(tag, q).encode(), mfd.value, -1), "b200_fd_send(mc)")
                lib.b200_fd_close(mfd.value)
            else:
                rfd = early_mc_fd
                if rfd is None:
                    src = C.c_int()
                    rfd = lib.b200_fd_recv(sock, C.byref(src))
                    if rfd < 0 or src.value != -1:
                        raise nv.NativeError("expected the multicast descriptor from rank 0: " + lib.b200_last_error().decode())
                nv._check(lib.b200_mc_import(rfd, C.byref(mc)), "b200_mc_import")
                lib.b200_fd_close(rfd)
            nv._check(lib.b200_mc_attach(mc.value, 1, 0, 0, None), "b200_mc_attach(add device)")
            dist.barrier(group=group)                      # every device must be added before anyone binds
            mp = C.c_void_p()
            nv._check(lib.b200_mc_attach(mc.value, 2, self.handle, self.nbytes, C.byref(mp)), "b200_mc_attach(bind)")
            dist.barrier(group=group)
            self.mc_ptr, self.mc_handle = mp.value, mc.value
        lib.b200_fd_close(sock)
        try:
            os.unlink(_sock_path(tag, rank))
        except OSError:
            pass

        from .symm import _CudaArray
        self._base = torch.as_tensor(_CudaArray(self.ptr, self.nbytes), device=self.device)
        self._cursor = 0
        self._base.zero_()                                 # cuMemCreate memory is not zeroed; flags and counters rely on it
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier(group=group)

    # ---- the SymmBuffer interface (parallel/symm.py) ---------------------------------------------------------------
    def tensor(self, offset_bytes: int, shape, dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        nb = n * torch.empty((), dtype=dtype).element_size()
        assert offset_bytes % 16 == 0 and offset_bytes + nb <= self.nbytes, (offset_bytes, nb, self.nbytes)
        return self._base[offset_bytes: offset_bytes + nb].view(dtype).view(*shape)

    def carve(self, shape, dtype: torch.dtype, align: int = 256):
        off = (self._cursor + align - 1) // align * align
        t = self.tensor(off, shape, dtype)
        self._cursor = off + t.numel() * t.element_size()
        return t, off

    def peers_at(self, offset_bytes: int):
        return (C.c_void_p * self.world)(*[p + offset_bytes for p in self.peer_ptrs])

    def mc_at(self, offset_bytes: int) -> Optional[int]:
        """The multicast address of ``offset_bytes`` inside the block (None without NVLS)."""
        return None if self.mc_ptr is None else self.mc_ptr + offset_bytes

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.ptr
        assert 0 <= off < self.nbytes, "tensor is not inside this symmetric block"
        return off

    def contains(self, t: torch.Tensor) -> bool:
        return self.ptr <= t.data_ptr() < self.ptr + self.nbytes

    def close(self) -> None:
        """Unmap the peers' blocks and free mine.  (The multicast object and its mapping are left to process exit: every
        rank would have to unbind before anyone releases, and nothing here reuses the address range.)"""
        lib = nv.load()
        self._base = None
        for q in range(self.world):
            if q != self.rank and self.peer_ptrs[q]:
                lib.b200_vmm_free(C.c_void_p(self.peer_ptrs[q]), self.nbytes, C.c_ulonglong(self.peer_handles[q]))
                self.peer_ptrs[q] = 0
        if self.ptr and self.mc_ptr is None:
            lib.b200_vmm_free(C.c_void_p(self.ptr), self.nbytes, C.c_ulonglong(self.handle))
        self.ptr = 0
