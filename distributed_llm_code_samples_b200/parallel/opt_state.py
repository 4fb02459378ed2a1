"""Optimizer state in the checkpoint's FULL, strategy-independent layout.

``{"kind": "adam", "t": steps, "m": [[...]], "v": [[...]]}`` with ``m`` / ``v`` shaped exactly like the full parameters
(fp32, CPU), so a run that stopped under one strategy resumes under any other with its Adam moments intact - the same
resharding rules as the parameters (FSDP: dim 0, ``train_ffns.py:266``; MP: W1/b1 dim 0, W2 dim 1, ``:317-319``).  The
reference has neither an optimizer state (plain SGD, ``:172``) nor checkpoints.

Every rank copies the pieces it owns into shared CPU tensors the parent allocated (``export_state``), and reads the pieces
it needs from them (``import_state``).  Trainers built on ``base.Optimizer`` are handled here; the fused DDP / FSDP
trainers, whose moments live in a compact per-owned-tile layout inside the kernel's channel, implement
``export_opt_state`` / ``import_opt_state`` themselves (``parallel/fused.py``).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from .base import tp_shard_dim


def owned_view(full: torch.Tensor, method: str, j: int, rank: int, world: int, for_write: bool) -> Optional[torch.Tensor]:
    """The slice of a full-layout CPU tensor that ``rank`` holds for parameter ``j`` of a layer under ``method``;
    ``None`` when another rank is responsible for writing a replicated tensor."""
    if method in ("single", "ddp"):
        return full if (rank == 0 or not for_write) else None
    if method == "fsdp":
        return full.chunk(world, dim=0)[rank]
    dim = tp_shard_dim(j)
    if dim is None:
        return full if (rank == 0 or not for_write) else None
    return full.chunk(world, dim=dim)[rank]


def _param_lists(method: str, trainer) -> Sequence:
    return trainer.params if method in ("single", "ddp") else trainer.shards


def _key(trainer, i: int, j: int, p: torch.Tensor):
    fn = getattr(trainer, "opt_key", None)
    return fn(i, j, p) if fn is not None else i * 4 + j


def export_state(method: str, trainer, comm, m_out: List[List[torch.Tensor]], v_out: List[List[torch.Tensor]]) -> int:
    """Write this rank's Adam moments into the shared full-layout tensors; returns the optimizer's step count."""
    if hasattr(trainer, "export_opt_state"):
        return trainer.export_opt_state(m_out, v_out)
    opt = getattr(trainer, "opt", None)
    if opt is None:
        return 0
    r, n = comm.rank, comm.world
    for i, lp in enumerate(_param_lists(method, trainer)):
        for j, p in enumerate(lp):
            st = opt.state.get(_key(trainer, i, j, p))
            if st is None:
                continue
            for src, dst_full in ((st[0], m_out[i][j]), (st[1], v_out[i][j])):
                dst = owned_view(dst_full, method, j, r, n, for_write=True)
                if dst is not None:
                    dst.copy_(src.detach().to("cpu", dtype=dst.dtype).reshape(dst.shape))
    return int(opt.t)


def import_state(method: str, trainer, comm, m_in: List[List[torch.Tensor]], v_in: List[List[torch.Tensor]], t: int) -> None:
    """Load this rank's pieces of a full-layout Adam state (no-op for SGD or ``t == 0``)."""
    if t <= 0:
        return
    if hasattr(trainer, "import_opt_state"):
        trainer.import_opt_state(m_in, v_in, t)
        return
    opt = getattr(trainer, "opt", None)
    if opt is None or opt.kind != "adam":
        return
    r, n = comm.rank, comm.world
    opt.t = int(t)
    for i, lp in enumerate(_param_lists(method, trainer)):
        for j, p in enumerate(lp):
            m = owned_view(m_in[i][j], method, j, r, n, for_write=False)
            v = owned_view(v_in[i][j], method, j, r, n, for_write=False)
            opt.state[_key(trainer, i, j, p)] = (m.to(device=p.device, dtype=torch.float32).reshape(p.shape).contiguous().clone(),
                                                 v.to(device=p.device, dtype=torch.float32).reshape(p.shape).contiguous().clone())


def empty_like_params(layers_params: Sequence[Sequence[torch.Tensor]]) -> List[List[torch.Tensor]]:
    return [[torch.zeros(p.shape, dtype=torch.float32).share_memory_() for p in lp] for lp in layers_params]
