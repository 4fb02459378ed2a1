"""Fully-sharded data-parallel trainer (reference ``train_process_fsdp``, ``train_ffns.py:197-259``).

Every parameter is sharded along dim 0.  A layer's full parameters are all-gathered one layer ahead of their use
in both passes (``:236,240,245,249``) into two pre-allocated, contiguous staging sets (the reference allocates
``n`` zero tensors per gather and concatenates, ``:202,:209``); the last layer's gathered parameters are reused
at the start of backward (``:244-246``).  Each layer's gradients are reduce-scattered (SUM) and applied to the
shard (``:253-259``) - asynchronously here, so the reduce-scatter of layer ``i`` overlaps the backward of layer
``i-1`` (the reference's open TODO, ``:14``, ``:252``).
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple

import torch

from ..config import TrainConfig
from ..models.ffn import Params, tlayer_ffn_bkwd, tlayer_ffn_fwd
from .base import Optimizer, block_opts


class FSDPTrainer:
    def __init__(self, ops, comm, cfg: TrainConfig, shards: List[Params]):
        self.ops, self.comm, self.cfg, self.shards = ops, comm, cfg, shards
        self.opts = block_opts(cfg)
        self.opt = Optimizer(ops, cfg)
        self.world = comm.world
        self.L = len(shards)
        # two staging sets of full parameters (double buffer for the one-layer-ahead prefetch)
        self.full = [[torch.empty((s.shape[0] * self.world, *s.shape[1:]), dtype=s.dtype, device=s.device)
                      for s in shards[0]] for _ in range(2)]
        self._pending = {}

    # -- parameter all-gather with one-layer-ahead prefetch ------------------------------------------------
    def _start_gather(self, layer: int) -> None:
        if not (0 <= layer < self.L) or layer in self._pending:
            return
        slot = layer % 2
        hs = [self.comm.all_gather_into(self.full[slot][k], self.shards[layer][k], async_op=True)
              for k in range(len(self.shards[layer]))]
        self._pending[layer] = hs

    def _finish_gather(self, layer: int) -> Params:
        for h in self._pending.pop(layer):
            h.wait()
        return self.full[layer % 2]

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        self.opt.begin_step()
        L = self.L
        acts = []
        y = x
        self._start_gather(0)
        params = None
        for l in range(L):                                       # forward (:238-241)
            params = self._finish_gather(l)
            self._start_gather(l + 1)
            y, s = tlayer_ffn_fwd(self.ops, params, y, self.opts)
            acts.append(s)
        self._start_gather(L - 2)                                # backward (:245); layer L-1 is still resident
        d = dloss_dx
        rs_prev: Optional[tuple] = None
        for i in reversed(range(L)):
            if i < L - 1:
                params = self._finish_gather(i)
                self._start_gather(i - 1)
            d, grads = tlayer_ffn_bkwd(self.ops, d, params, acts[i], self.opts, need_dx=i > 0)
            acts[i] = None
            if rs_prev is not None:                              # apply layer i+1's update while layer i ran
                self._apply(*rs_prev)
            # reduce-scatter outputs (:227) in the GRADIENT's dtype: bias gradients are fp32 sums whatever the parameters are
            outs = [torch.empty(s.shape, dtype=g.dtype, device=s.device) for s, g in zip(self.shards[i], grads)]
            hs = [self.comm.reduce_scatter_into(o, g, async_op=True) for o, g in zip(outs, grads)]
            rs_prev = (i, outs, hs, grads)
        self._apply(*rs_prev)
        return y

    def _apply(self, layer: int, outs, handles, grads_keepalive) -> None:
        for h in handles:
            h.wait()
        for j, (p, g) in enumerate(zip(self.shards[layer], outs)):
            self.opt.update(layer * 4 + j, p, g)                 # sharded update (:258-259)

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        """All-gather every shard (dim 0) - what the reference's parent does with ``torch.cat`` (``:284-287``)."""
        out = []
        for lp in self.shards:
            fl = []
            for s in lp:
                full = torch.empty((s.shape[0] * self.world, *s.shape[1:]), dtype=s.dtype, device=s.device)
                self.comm.all_gather_into(full, s)
                fl.append(full)
            out.append(fl)
        return out
