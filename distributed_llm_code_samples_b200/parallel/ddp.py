"""Data-parallel trainer (reference ``train_process_ddp``, ``train_ffns.py:156-172``).

Parameters are replicated, every rank sees its own batches, and each layer's gradients are all-reduced (SUM,
no averaging) as soon as that layer's backward has produced them, overlapping with the backward of the layers
below (``:90-91``, ``:164-165``); the updates happen after the waits, last layer first (``:168-172``).
"""
from __future__ import annotations

from typing import Iterable, List, Tuple

import torch

from ..config import TrainConfig
from ..models.ffn import Params, tlayers_ffn_bkwd, tlayers_ffn_fwd
from .base import Optimizer, block_opts


class DDPTrainer:
    def __init__(self, ops, comm, cfg: TrainConfig, params: List[Params]):
        self.ops, self.comm, self.cfg, self.params = ops, comm, cfg, params
        self.opts = block_opts(cfg)
        self.opt = Optimizer(ops, cfg)

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        self.opt.begin_step()
        y, acts = tlayers_ffn_fwd(self.ops, self.params, x, self.opts)

        def ddp_comms_hook(i, grads):
            return [self.comm.all_reduce(g, async_op=True) for g in grads]

        grads, handles = tlayers_ffn_bkwd(self.ops, dloss_dx, self.params, acts, self.opts, after_comms_hook=ddp_comms_hook)
        for i in reversed(range(len(self.params))):
            for h in handles[i]:
                h.wait()
            for j, (p, g) in enumerate(zip(self.params[i], grads[i])):
                self.opt.update(i * 4 + j, p, g)
        return y

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        return self.params
