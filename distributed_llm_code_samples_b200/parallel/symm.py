"""Symmetric memory and the native (peer-memory) collectives.

``SymmWorld`` owns the per-rank side of the peer table: every rank allocates the same-sized block with the native
allocator, the CUDA-IPC handles are exchanged once over ``torch.distributed`` (bootstrap only), and every peer's
block is mapped.  This is the explicit version of what the reference gets by pickling CUDA tensors into child
processes (``train_ffns.py:186``; ``test_mp_barrier_gpus.py:33``).

``SymmComm`` implements the strategies' collective interface (parallel/comm.py) on the native kernels of
``csrc/comm.cu`` - loads/stores on NVLink-mapped peer pointers, device-side flag barriers, no NCCL on the data path.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native as nv


class _CudaArray:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class SymmBuffer:
    """One symmetric allocation: ``ptr`` (own), ``peer_ptrs`` (all ranks, own at index rank)."""

    def __init__(self, world: "SymmWorld", nbytes: int):
        lib = nv.load()
        self.world, self.nbytes = world, nbytes
        p = C.c_void_p()
        nv._check(lib.b200_symm_alloc(C.byref(p), nbytes), "b200_symm_alloc")
        self.ptr = p.value
        self._opened: List[int] = []
        if world.world > 1:
            hb = lib.b200_ipc_handle_bytes()
            h = (C.c_ubyte * hb)()
            nv._check(lib.b200_ipc_get_handle(h, C.c_void_p(self.ptr)), "b200_ipc_get_handle")
            handles: List[Optional[bytes]] = [None] * world.world
            dist.all_gather_object(handles, bytes(h), group=world.group)
            ptrs = []
            for r, hr in enumerate(handles):
                if r == world.rank:
                    ptrs.append(self.ptr)
                    continue
                q = C.c_void_p()
                buf = (C.c_ubyte * hb).from_buffer_copy(hr)
                nv._check(lib.b200_ipc_open_handle(C.byref(q), buf), "b200_ipc_open_handle")
                ptrs.append(q.value)
                self._opened.append(q.value)
            self.peer_ptrs = ptrs
        else:
            self.peer_ptrs = [self.ptr]
        self.peers_c = (C.c_void_p * world.world)(*self.peer_ptrs)
        self.mc_ptr: Optional[int] = None          # (VmmBuffer blocks have one where the box does NVLS)
        self._base = torch.as_tensor(_CudaArray(self.ptr, nbytes), device=world.device)
        self._cursor = 0

    def tensor(self, offset_bytes: int, shape, dtype: torch.dtype) -> torch.Tensor:
        """Zero-copy torch view of OWN memory at ``offset_bytes``."""
        n = 1
        for s in shape:
            n *= s
        nb = n * torch.empty((), dtype=dtype).element_size()
        assert offset_bytes % 16 == 0 and offset_bytes + nb <= self.nbytes, (offset_bytes, nb, self.nbytes)
        return self._base[offset_bytes: offset_bytes + nb].view(dtype).view(*shape)

    def carve(self, shape, dtype: torch.dtype, align: int = 256):
        """Bump-allocate a tensor inside the block; returns ``(tensor, offset_bytes)`` (same offsets on every rank)."""
        off = (self._cursor + align - 1) // align * align
        t = self.tensor(off, shape, dtype)
        self._cursor = off + t.numel() * t.element_size()
        return t, off

    def peers_at(self, offset_bytes: int):
        """ctypes array of every rank's pointer to ``offset_bytes`` inside this block."""
        return (C.c_void_p * self.world.world)(*[p + offset_bytes for p in self.peer_ptrs])

    def mc_at(self, offset_bytes: int) -> Optional[int]:
        return None

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.ptr
        assert 0 <= off < self.nbytes, "tensor is not inside this symmetric block"
        return off

    def contains(self, t: torch.Tensor) -> bool:
        return self.ptr <= t.data_ptr() < self.ptr + self.nbytes

    def close(self) -> None:
        lib = nv.load()
        for q in self._opened:
            lib.b200_ipc_close_handle(C.c_void_p(q))
        self._opened = []
        if self.ptr:
            self._base = None
            lib.b200_symm_free(C.c_void_p(self.ptr))
            self.ptr = 0


class SymmWorld:
    """Per-rank handle on the symmetric-memory world (bootstrap over an existing process group)."""

    def __init__(self, rank: int, world: int, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        self.rank, self.world, self.device, self.group = rank, world, device, group
        self.buffers: List[SymmBuffer] = []
        self._mc_ok: Optional[bool] = None
        self._vmm_count = 0
        torch.cuda.set_device(device)

    def multicast_supported(self) -> bool:
        """NVLS on this box: POSIX-fd shareable cuMem blocks + multicast objects, and more than one rank to bind."""
        if self._mc_ok is None:
            from . import vmm
            fd_ok, mc_ok = vmm.supported() if self.world > 1 else (False, False)
            ok = torch.tensor([int(fd_ok and mc_ok)], device=self.device)
            if self.world > 1:
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)     # all ranks or none
            self._mc_ok = bool(ok.item())
        return self._mc_ok

    def alloc(self, nbytes: int, multicast: bool = False):
        """Collective: every rank must call with the same size in the same order.  ``multicast=True`` (where the box has
        NVLS) returns a cuMem VMM block bound to a multicast object (``.mc_ptr``), else a cudaMalloc + CUDA-IPC block."""
        nbytes = (nbytes + 255) // 256 * 256
        b = None
        if multicast and self.multicast_supported():
            from .vmm import VmmBuffer
            self._vmm_count += 1
            err = None
            try:
                b = VmmBuffer(self.rank, self.world, self.device, nbytes, multicast=True, group=self.group, tag=f"sw{self._vmm_count}_")
            except Exception as e:  # noqa: BLE001 - a box without working NVLS falls back to IPC blocks, on EVERY rank
                err = e
            ok = torch.tensor([0 if (err is not None or b is None or not b.mc_ptr) else 1], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
            if not bool(ok.item()):
                import warnings
                warnings.warn(f"NVLS multicast block unavailable ({err}); falling back to CUDA-IPC symmetric memory")
                self._mc_ok = False
                b = None
        if b is None:
            b = SymmBuffer(self, nbytes)
        self.buffers.append(b)
        return b

    def host_barrier(self) -> None:
        if self.world > 1:
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)

    def close(self) -> None:
        self.host_barrier()
        for b in self.buffers:
            b.close()
        self.buffers = []


class _Handle:
    def __init__(self, event: Optional[torch.cuda.Event]):
        self.event = event

    def wait(self) -> None:
        if self.event is not None:
            torch.cuda.current_stream().wait_event(self.event)


class CollChannel:
    """Flag block + epoch counter for one stream of native collectives."""

    def __init__(self, sw: SymmWorld):
        self.sw = sw
        self.flags = sw.alloc(int(nv.load().b200_coll_flag_bytes()))
        self.epoch = 0

    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch


class SymmComm:
    """The strategies' collective interface on the native peer-memory kernels (bf16 payloads).

    Operands that already live in symmetric memory are used in place; others are staged through a symmetric
    bounce buffer (one local copy in, one out).

    Two independent lanes: blocking calls run on the caller's current stream, ``async_op=True`` calls on ``self.side``.
    Each lane has its OWN flag block, epoch counter and half of the staging buffer, so a caller may mix the two forms
    freely (kernels of different lanes can overtake each other on the device without ever seeing each other's flags or
    staging bytes).  Within a lane the order is the stream's.  All blocking calls must be issued on the same stream."""

    name = "b200-p2p"

    def __init__(self, sw: SymmWorld, staging_bytes: int = 64 << 20, group: Optional[dist.ProcessGroup] = None):
        self.sw, self.rank, self.world, self.group = sw, sw.rank, sw.world, group
        self.chans = (CollChannel(sw), CollChannel(sw))          # [0] blocking lane, [1] async (side-stream) lane
        self.chan = self.chans[0]
        half = (staging_bytes + 255) // 256 * 256
        self.staging = sw.alloc(2 * half, multicast=True)       # NVLS-capable where the box has it (else a plain IPC block)
        self._stage_half = half
        self.side = torch.cuda.Stream(device=sw.device)
        self.lib = nv.load()

    # ---- helpers ---------------------------------------------------------------------------------------------
    def _run(self, fn, async_op: bool):
        """Run ``fn(raw_stream_handle)`` with torch's current stream set to that same stream, so the torch copies
        and the native kernel inside ``fn`` are ordered with each other."""
        if not async_op:
            fn(torch.cuda.current_stream().cuda_stream)
            nv.launch_count += 1
            return None
        self.side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            fn(self.side.cuda_stream)
            ev = torch.cuda.Event()
            ev.record(self.side)
        nv.launch_count += 1
        return _Handle(ev)

    def _find(self, t: torch.Tensor) -> Optional[SymmBuffer]:
        for b in self.sw.buffers:
            if b.contains(t):
                return b
        return None

    def _lane(self, async_op: bool):
        """(channel, staging offset) of the lane a call runs in."""
        return (self.chans[1], self._stage_half) if async_op else (self.chans[0], 0)

    def _reduce(self, op: int, buf, off: int, nbytes: int, f32: bool, chan: "CollChannel", e: int, s, fallback) -> None:
        """All-reduce (op 0) / reduce-scatter (op 1) of ``nbytes`` at ``off`` of symmetric block ``buf``: inside the
        NVSwitch when the block is multicast-bound (multimem.ld_reduce / multimem.st), else on the peer-pointer kernels."""
        # Two ranks gain nothing from the switch on large messages (measured, profiles/bench/r2/coll_sweep_n2_nvls.jsonl: the
        # peer-pointer two-shot moves the same bytes with one barrier less); from four ranks on NVLS halves the traffic.
        use_nvls = getattr(buf, "mc_ptr", None) and (self.world >= 4 or nbytes <= (512 << 10))
        if use_nvls:
            nv._check(self.lib.b200_nvls_collective(op, int(f32), buf.mc_ptr + off, buf.ptr + off, nbytes, chan.flags.peers_c,
                                                    self.rank, self.world, e, 0, s), "b200_nvls_collective")
        else:
            nv._check(fallback(buf.peers_c, off, nbytes, chan.flags.peers_c, self.rank, self.world, e, 0, s),
                      "b200_all_reduce" if op == 0 else "b200_reduce_scatter")

    def _stage_check(self, nbytes: int) -> None:
        if nbytes > self._stage_half:
            raise ValueError(f"operand of {nbytes} bytes exceeds the symmetric staging buffer ({self._stage_half}); "
                             "allocate it with SymmWorld.alloc or raise staging_bytes")

    # ---- collectives ----------------------------------------------------------------------------------------
    def barrier(self) -> None:
        e = self.chan.next_epoch()
        nv._check(self.lib.b200_barrier(self.chan.flags.peers_c, self.rank, self.world, e, nv._stream()), "b200_barrier")
        nv.launch_count += 1

    def all_reduce(self, t: torch.Tensor, async_op: bool = False):
        assert t.dtype in (torch.bfloat16, torch.float32) and t.is_contiguous(), "native all_reduce: contiguous bf16 / fp32"
        es = t.element_size()
        fn = self.lib.b200_all_reduce_bf16 if es == 2 else self.lib.b200_all_reduce_f32
        nbytes = (t.numel() * es + 15) // 16 * 16
        buf = self._find(t)
        chan, soff = self._lane(async_op)
        e = chan.next_epoch()
        if buf is not None and (t.numel() * es) % 16 == 0:
            off = buf.offset_of(t)
            return self._run(lambda s: self._reduce(0, buf, off, nbytes, es == 4, chan, e, s, fn), async_op)
        self._stage_check(nbytes)
        st = self.staging.tensor(soff, (nbytes // es,), t.dtype)

        def go(s):
            st[t.numel():].zero_()                                   # the padding lanes are summed too
            st[: t.numel()].copy_(t.reshape(-1))
            self._reduce(0, self.staging, soff, nbytes, es == 4, chan, e, s, fn)
            t.reshape(-1).copy_(st[: t.numel()])
        return self._run(go, async_op)

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor, async_op: bool = False):
        assert out.is_contiguous() and shard.is_contiguous()
        sb = shard.numel() * shard.element_size()
        assert sb % 16 == 0, "shard size must be a multiple of 16 bytes"
        chan, soff = self._lane(async_op)
        e = chan.next_epoch()
        buf = self._find(shard)
        if buf is not None:
            off = buf.offset_of(shard)
            return self._run(lambda s: nv._check(self.lib.b200_all_gather_into(out.data_ptr(), buf.peers_c, off, sb, chan.flags.peers_c,
                                                                              self.rank, self.world, e, 0, s), "b200_all_gather_into"), async_op)
        self._stage_check(sb)
        st = self.staging.tensor(soff, (sb,), torch.uint8)

        def go(s):
            st.copy_(shard.reshape(-1).view(torch.uint8))
            nv._check(self.lib.b200_all_gather_into(out.data_ptr(), self.staging.peers_c, soff, sb, chan.flags.peers_c,
                                                    self.rank, self.world, e, 0, s), "b200_all_gather_into")
        return self._run(go, async_op)

    def reduce_scatter_into(self, out: torch.Tensor, full: torch.Tensor, async_op: bool = False):
        assert full.dtype in (torch.bfloat16, torch.float32) and full.is_contiguous() and out.is_contiguous()
        es = full.element_size()
        rs = self.lib.b200_reduce_scatter_bf16 if es == 2 else self.lib.b200_reduce_scatter_f32
        nbytes = full.numel() * es
        assert nbytes % (16 * self.world) == 0, "reduce_scatter needs bytes % (16*world) == 0"
        chan, soff = self._lane(async_op)
        e = chan.next_epoch()
        buf = self._find(full)
        sl = nbytes // self.world

        if buf is not None:
            off = buf.offset_of(full)

            def go(s):
                self._reduce(1, buf, off, nbytes, es == 4, chan, e, s, rs)
                out.reshape(-1).copy_(full.reshape(-1)[self.rank * (sl // es): (self.rank + 1) * (sl // es)])
            return self._run(go, async_op)
        self._stage_check(nbytes)
        st = self.staging.tensor(soff, (nbytes // es,), full.dtype)

        def go2(s):
            st.copy_(full.reshape(-1))
            self._reduce(1, self.staging, soff, nbytes, es == 4, chan, e, s, rs)
            out.reshape(-1).copy_(st[self.rank * (sl // es): (self.rank + 1) * (sl // es)])
        return self._run(go2, async_op)

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)     # control plane, not the data path
        return t


_WORLD: Optional[SymmWorld] = None


def get_symm_world(comm, device: torch.device) -> SymmWorld:
    """Process-wide symmetric world for the default group (created on first use)."""
    global _WORLD
    if _WORLD is None:
        _WORLD = SymmWorld(comm.rank, comm.world, device, getattr(comm, "group", None))
    return _WORLD


def release_world() -> None:
    """Free every symmetric block of the process-wide world (collective: every rank calls it).  Tensors carved out of
    the blocks must already be unreachable.  The next ``get_symm_world`` starts a fresh world."""
    global _WORLD
    if _WORLD is not None:
        _WORLD.close()
        _WORLD = None
