"""Collective layer used by the strategies.

``TorchComm`` wraps ``torch.distributed`` (NCCL on GPUs - the reference's transport, ``train_ffns.py:165,203,
255,303`` - or gloo on CPU for the plumbing tests).  ``SymmComm`` (parallel/symm.py) provides the same interface
on the framework's own peer-memory kernels.  Tensor-form collectives are used throughout: a dim-0 shard is a
contiguous slice of the full tensor, so neither the reference's per-call ``zeros_like`` lists (``:202``) nor its
``torch.cat`` (``:209``) are needed.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


class _Done:
    def wait(self) -> None:
        return None


class LocalComm:
    """World of one: every collective is the identity."""
    rank, world = 0, 1
    name = "local"

    def all_reduce(self, t: torch.Tensor, async_op: bool = False):
        return _Done() if async_op else None

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor, async_op: bool = False):
        out.view(-1)[: shard.numel()].copy_(shard.reshape(-1))
        return _Done() if async_op else None

    def reduce_scatter_into(self, out: torch.Tensor, full: torch.Tensor, async_op: bool = False):
        out.copy_(full.view(-1)[: out.numel()].view_as(out))
        return _Done() if async_op else None

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def barrier(self) -> None:
        return None


class TorchComm:
    name = "torch.distributed"

    def __init__(self, group: Optional[dist.ProcessGroup] = None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.backend = dist.get_backend(group)

    def all_reduce(self, t: torch.Tensor, async_op: bool = False):
        return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor, async_op: bool = False):
        """``out`` (world x shard, contiguous) <- concatenation of every rank's ``shard`` along dim 0."""
        return dist.all_gather_into_tensor(out, shard.contiguous(), group=self.group, async_op=async_op)

    def reduce_scatter_into(self, out: torch.Tensor, full: torch.Tensor, async_op: bool = False):
        """``out`` <- this rank's dim-0 chunk of the element-wise SUM of ``full`` over ranks."""
        if self.backend == "gloo":
            # gloo has no reduce_scatter: all-reduce a copy and keep our slice (plumbing path only)
            tmp = full.clone()
            dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=self.group)
            out.copy_(tmp.chunk(self.world, dim=0)[self.rank])
            return _Done() if async_op else None
        return dist.reduce_scatter_tensor(out, full.contiguous(), op=dist.ReduceOp.SUM, group=self.group,
                                          async_op=async_op)

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t

    def barrier(self) -> None:
        dist.barrier(group=self.group)


def make_comm(group: Optional[dist.ProcessGroup] = None):
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        return TorchComm(group)
    return LocalComm()
