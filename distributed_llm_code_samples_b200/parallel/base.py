"""Shared pieces of the per-rank trainers: optimizer state, block options, sharding rules."""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..config import TrainConfig
from ..models.ffn import BlockOpts, Params


def block_opts(cfg: TrainConfig) -> BlockOpts:
    return BlockOpts(activation=cfg.activation, residual=cfg.residual, recompute=cfg.recompute)


def set_update_rounding(cfg: TrainConfig, ops, step: int) -> None:
    """Select the rounding of this step's bf16 weight updates on the native kernels (``b200_set_update_rounding``): a fresh
    seed per step when ``cfg.stochastic_rounding``, round-to-nearest otherwise.  Same seed on every rank."""
    if getattr(ops, "name", "") != "native":
        return
    from ..ops import native as nv
    nv.set_update_rounding(((0x9E3779B1 * (step + 1)) & 0xFFFFFFFF) | 1 if cfg.stochastic_rounding else 0)


class Optimizer:
    """SGD (the reference's ``param.add_(-LR*grad)``, ``train_ffns.py:172``) or Adam with fp32 moments.
    ``grad_scale`` multiplies the summed gradient (reference: 1.0 = plain SUM over ranks, ``:165``)."""

    def __init__(self, ops, cfg: TrainConfig):
        self.ops, self.kind, self.cfg = ops, cfg.optimizer, cfg
        self.lr = cfg.lr * cfg.grad_scale
        self.beta1, self.beta2, self.eps = 0.9, 0.999, 1e-8
        self.state: Dict[int, Tuple[torch.Tensor, torch.Tensor]] = {}
        self.t = 0

    def begin_step(self) -> None:
        self.t += 1
        set_update_rounding(self.cfg, self.ops, self.t)

    def update(self, key: int, p: torch.Tensor, g: Optional[torch.Tensor]) -> None:
        if g is None:                      # already applied inside the GEMM that produced it
            return
        if self.kind == "sgd":
            self.ops.sgd_(p, g, self.lr)
            return
        if key not in self.state:
            self.state[key] = (torch.zeros_like(p, dtype=torch.float32), torch.zeros_like(p, dtype=torch.float32))
        m, v = self.state[key]
        self.ops.adam_(p, g, m, v, self.lr, self.beta1, self.beta2, self.eps, self.t)

    def state_dict(self) -> dict:
        return {"kind": self.kind, "t": self.t, "state": {k: (m.cpu(), v.cpu()) for k, (m, v) in self.state.items()}}

    def load_state_dict(self, sd: dict, device: torch.device) -> None:
        self.t = sd.get("t", 0)
        self.state = {k: (m.to(device), v.to(device)) for k, (m, v) in sd.get("state", {}).items()}


# ------------------------------------------------------------------------------------------- sharding rules
def fsdp_shard(p: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """FSDP shards EVERY parameter along dim 0 (``train_ffns.py:265-266``)."""
    if p.shape[0] % world:
        raise ValueError(f"FSDP needs dim 0 ({p.shape[0]}) divisible by the world size ({world})")
    return p.chunk(world, dim=0)[rank]


def tp_shard_dim(index_in_layer: int) -> Optional[int]:
    """Megatron MP: W1 (index 0) by rows = column-parallel, W2 (index 1) by columns = row-parallel
    (``train_ffns.py:316-319``); b1 (2) follows W1; b2 (3) is replicated."""
    return {0: 0, 1: 1, 2: 0, 3: None}[index_in_layer]


def tp_shard(p: torch.Tensor, index_in_layer: int, rank: int, world: int) -> torch.Tensor:
    dim = tp_shard_dim(index_in_layer)
    if dim is None:
        return p
    if p.shape[dim] % world:
        raise ValueError(f"MP needs dim {dim} ({p.shape[dim]}) divisible by the world size ({world})")
    return p.chunk(world, dim=dim)[rank]


def place(layers_params: Sequence[Params], device: torch.device, dtype: torch.dtype) -> List[Params]:
    # copy=True: on CPU the inputs are the parent's shared-memory tensors; every rank needs a PRIVATE replica
    # biases (index 2, 3) stay fp32: the GEMM epilogues take fp32 bias vectors and the bias gradients are fp32 sums
    return [[p.to(device=device, dtype=dtype if j < 2 else torch.float32, copy=True).contiguous() for j, p in enumerate(lp)]
            for lp in layers_params]
