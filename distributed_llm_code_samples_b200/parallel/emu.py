"""Several ranks of the fused GEMM+collective kernels on ONE GPU, as slices of one cooperative launch.

Kernels that wait on one another's flags cannot be run as separate launches on one device (nothing makes them
co-resident; the GPU box's guide forbids it), so the native library offers ``b200_emu_*`` entry points in which CTAs
``[r * cpr, (r + 1) * cpr)`` of a single grid execute rank ``r``'s kernel body - its own operands, owner maps, tile walk,
flags - against peer tables whose every entry is local memory (``csrc/gemm_kernel.cuh``: ``gemm_tcgen05_emu_kernel``).
This is how the cross-rank protocols of ``csrc/fused.cu`` and ``csrc/gather_gemm.cu`` are tested (and profiled with ncu)
on a single B200; the multi-GPU runs then only add NVLink.

``EmuFused`` / ``EmuGather`` mirror ``FusedChannel`` / ``GatherChannel`` of ``parallel/fused.py`` with per-rank lists in
place of per-rank processes.  No counterpart in the reference (its collectives are NCCL calls, ``train_ffns.py:165,203,255``).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from ..ops import native as nv

OWN_INTERLEAVED, OWN_ROW_SHARDED = 0, 1
FOP_SUM, FOP_SGD = 0, 1


def _tab(ptrs: Sequence[int]):
    return (C.c_void_p * len(ptrs))(*ptrs)


def ctas_per_rank(world: int) -> int:
    """Largest even slice such that ``world`` slices fit the device's SMs (one CTA per SM)."""
    return max(2, (nv.num_sms() // world) // 2 * 2)


class EmuFused:
    """Inbox + flags of ``world`` emulated ranks and the per-channel call counter (cf. ``FusedChannel``)."""

    def __init__(self, world: int, owned_cap: int, ready_cap: int, device: torch.device):
        lib = nv.load()
        self.world, self.owned_cap, self.ready_cap, self.device = world, owned_cap, ready_cap, device
        ib, fb = int(lib.b200_fused_inbox_bytes(owned_cap, world)), int(lib.b200_fused_flag_bytes(owned_cap, ready_cap))
        self.inbox = [torch.zeros(ib, dtype=torch.uint8, device=device) for _ in range(world)]
        self.flags = [torch.zeros(fb, dtype=torch.uint8, device=device) for _ in range(world)]
        self.table = torch.zeros(int(lib.b200_emu_table_bytes()), dtype=torch.uint8, device=device)
        self.inbox_c, self.flags_c = _tab([t.data_ptr() for t in self.inbox]), _tab([t.data_ptr() for t in self.flags])
        self.seq = 0

    def gemm(self, A: List[torch.Tensor], a_major: int, B: List[torch.Tensor], b_major: int, M: int, N: int, K: int,
             out: List[torch.Tensor], owner_mode: int, op: int, broadcast: bool, alpha: float = 1.0,
             adam_state: Optional[List] = None, lr: float = 0.0, step: int = 1, beta1: float = 0.9, beta2: float = 0.999,
             eps: float = 1e-8, grad_scale: float = 1.0, cpr: int = 0) -> None:
        """Rank r computes ``A[r] . B[r]`` ; tile sums land in ``out`` (every rank's full ``[M, N]`` when ``broadcast``,
        else rank r's ``[M / world, N]`` row shard)."""
        self.seq += 1
        w = self.world
        am = _tab([s[0].data_ptr() for s in adam_state]) if adam_state is not None else None
        av = _tab([s[1].data_ptr() for s in adam_state]) if adam_state is not None else None
        rc = nv.load().b200_emu_gemm_allreduce_bf16(
            w, _tab([a.data_ptr() for a in A]), A[0].stride(0), a_major, _tab([b.data_ptr() for b in B]), B[0].stride(0), b_major,
            M, N, K, _tab([o.data_ptr() for o in out]), out[0].stride(0), self.inbox_c, self.flags_c, self.owned_cap, self.ready_cap,
            self.seq, owner_mode, op, int(broadcast), float(alpha), am, av, float(lr), float(beta1), float(beta2), float(eps),
            int(step), float(grad_scale), cpr or ctas_per_rank(w), self.table.data_ptr(), nv._stream())
        nv._check(rc, "b200_emu_gemm_allreduce_bf16")
        nv.launch_count += 1


class EmuGather:
    """``started`` flags and ``ready`` counters of ``world`` emulated ranks for the all-gather + GEMM kernel."""

    def __init__(self, world: int, max_rows: int, device: torch.device):
        lib = nv.load()
        self.world, self.device = world, device
        self.started = [torch.zeros(64, dtype=torch.int32, device=device) for _ in range(world)]
        self.ready = [torch.zeros(int(lib.b200_gather_gemm_ready_bytes(max_rows)) // 4, dtype=torch.int32, device=device) for _ in range(world)]
        self.table = torch.zeros(int(lib.b200_emu_table_bytes()), dtype=torch.uint8, device=device)
        self.seq = 0

    def gemm(self, a_bufs: List[torch.Tensor], a_major: int, B: List[torch.Tensor], b_major: int, out: List[torch.Tensor], M: int, N: int,
             K: int, epilogue: int, aux: Optional[List[torch.Tensor]] = None, out2: Optional[List[torch.Tensor]] = None, cpr: int = 0) -> None:
        """``a_bufs[r]``: rank r's full-size A buffer in which only its own shard is valid on entry (complete on return)."""
        self.seq += 1
        w = self.world
        rc = nv.load().b200_emu_gather_gemm_bf16(
            w, _tab([a.data_ptr() for a in a_bufs]), a_major, _tab([b.data_ptr() for b in B]), B[0].stride(0), b_major,
            _tab([o.data_ptr() for o in out]), out[0].stride(0), M, N, K, epilogue,
            _tab([t.data_ptr() for t in aux]) if aux is not None else None, aux[0].stride(0) if aux is not None else 0,
            _tab([t.data_ptr() for t in out2]) if out2 is not None else None, out2[0].stride(0) if out2 is not None else 0,
            _tab([t.data_ptr() for t in self.started]), _tab([t.data_ptr() for t in self.ready]), self.seq,
            cpr or ctas_per_rank(w), self.table.data_ptr(), nv._stream())
        nv._check(rc, "b200_emu_gather_gemm_bf16")
        nv.launch_count += 1
