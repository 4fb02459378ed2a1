"""Single-device trainer (reference ``train_1gpu``, ``train_ffns.py:101-116``)."""
from __future__ import annotations

from typing import Iterable, List, Tuple

import torch

from ..config import TrainConfig
from ..models.ffn import Params, tlayer_ffn_bkwd, tlayers_ffn_bkwd, tlayers_ffn_fwd
from .base import Optimizer, block_opts


class SingleTrainer:
    """fwd -> bwd -> SGD on one device.  With the native ops and SGD the update is fused into the dW GEMM's
    epilogue, so no gradient tensor is ever written (the reference builds new parameter tensors, ``:114``)."""

    def __init__(self, ops, cfg: TrainConfig, params: List[Params]):
        self.ops, self.cfg, self.params = ops, cfg, params
        self.opts = block_opts(cfg)
        self.opt = Optimizer(ops, cfg)
        self.fused = bool(getattr(ops, "fused_update", False)) and cfg.optimizer == "sgd"

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        self.opt.begin_step()
        y, acts = tlayers_ffn_fwd(self.ops, self.params, x, self.opts)
        if self.fused:
            d = dloss_dx
            for i in reversed(range(len(self.params))):
                d, g = tlayer_ffn_bkwd(self.ops, d, self.params[i], acts[i], self.opts, need_dx=i > 0,
                                       update_lr=self.opt.lr)
                acts[i] = None      # release this block's activations as soon as they are consumed
                if g is not None:   # bias gradients (the weights were updated in the dW GEMMs' epilogues)
                    for j, (p, gj) in enumerate(zip(self.params[i], g)):
                        self.opt.update(i * 4 + j, p, gj)
            return y
        grads, _ = tlayers_ffn_bkwd(self.ops, dloss_dx, self.params, acts, self.opts)
        for i in range(len(self.params)):
            for j, (p, g) in enumerate(zip(self.params[i], grads[i])):
                self.opt.update(i * 4 + j, p, g)
        return y

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        return self.params
