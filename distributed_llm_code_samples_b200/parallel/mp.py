"""Megatron-style model (tensor) parallel trainer (reference ``train_process_tp``, ``train_ffns.py:290-312``).

``W1`` is split by rows (column-parallel), ``W2`` by columns (row-parallel) (``:316-319``); every rank sees the
same batches (``:324``).  Two communication forms:

* ``ar`` (reference): activations replicated, ``all_reduce(y)`` after each block's forward (``:303``) and
  ``all_reduce(dx)`` after each block's backward (``:309``).
* ``sp`` (sequence-parallel extension): between blocks activations are sharded over tokens;
  block input is all-gathered, block output reduce-scattered.  Same traffic (AR = RS + AG), 1/n of the saved
  activation memory.

Weight gradients are shard-local: no gradient communication (``:311-312``).
"""
from __future__ import annotations

from typing import Iterable, List, Tuple

import torch

from ..config import TrainConfig
from ..models.ffn import Params, tlayer_ffn_bkwd, tlayer_ffn_fwd
from .base import Optimizer, block_opts


class MPTrainer:
    def __init__(self, ops, comm, cfg: TrainConfig, shards: List[Params]):
        self.ops, self.comm, self.cfg, self.shards = ops, comm, cfg, shards
        self.opts = block_opts(cfg)
        if cfg.residual:
            # the residual is added once, after the partial sums are reduced
            self.opts.residual = False
        self.residual = cfg.residual
        self.opt = Optimizer(ops, cfg)
        self.world, self.rank = comm.world, comm.rank
        self.fused = bool(getattr(ops, "fused_update", False)) and cfg.optimizer == "sgd"
        self.sp = cfg.mp_mode in ("sp", "spf")      # spf is a fused-kernel variant of sp: same math on this path

    def _params_for_fwd(self, p: Params) -> Params:
        # b2 must enter the reduced sum exactly once
        if len(p) == 4 and self.rank != 0:
            return [p[0], p[1], p[2], torch.zeros_like(p[3])]
        return p

    # ---- reference form ---------------------------------------------------------------------------------
    def _step_ar(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        acts = []
        y = x
        for p in self.shards:
            inp = y
            y, s = tlayer_ffn_fwd(self.ops, self._params_for_fwd(p), y, self.opts)
            self.comm.all_reduce(y)                                           # (:303)
            if self.residual:
                y = y + inp
            acts.append(s)
        d = dloss_dx
        for i in reversed(range(len(self.shards))):
            dn, grads = tlayer_ffn_bkwd(self.ops, d, self.shards[i], acts[i], self.opts, need_dx=i > 0,
                                        update_lr=self.opt.lr if self.fused else None)
            acts[i] = None
            if dn is not None:
                self.comm.all_reduce(dn)                                      # (:309)
                if self.residual:
                    dn = dn + d
            if grads is not None:
                self._update(i, grads)
            d = dn
        return y

    # ---- sequence-parallel form -------------------------------------------------------------------------
    def _step_sp(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        n, r = self.world, self.rank
        T = x.shape[0]
        if T % n:
            raise ValueError(f"sequence-parallel MP needs tokens ({T}) divisible by the world size ({n})")
        xs = x.chunk(n, dim=0)[r].contiguous()
        acts = []
        for p in self.shards:
            full = torch.empty((T, xs.shape[1]), dtype=xs.dtype, device=xs.device)
            self.comm.all_gather_into(full, xs)                               # AG -> column-parallel GEMM
            y, s = tlayer_ffn_fwd(self.ops, self._params_for_fwd(p), full, self.opts)
            s.x = xs                                                          # keep only the token shard
            out = torch.empty_like(xs)
            self.comm.reduce_scatter_into(out, y)                             # row-parallel GEMM -> RS
            if self.residual:
                out = out + xs
            acts.append(s)
            xs = out
        d = dloss_dx.chunk(n, dim=0)[r].contiguous()
        for i in reversed(range(len(self.shards))):
            dfull = torch.empty((T, d.shape[1]), dtype=d.dtype, device=d.device)
            self.comm.all_gather_into(dfull, d)
            s = acts[i]
            xfull = torch.empty((T, s.x.shape[1]), dtype=s.x.dtype, device=s.x.device)
            self.comm.all_gather_into(xfull, s.x)
            s.x = xfull
            dn, grads = tlayer_ffn_bkwd(self.ops, dfull, self.shards[i], s, self.opts, need_dx=i > 0,
                                        update_lr=self.opt.lr if self.fused else None)
            acts[i] = None
            if dn is not None:
                out = torch.empty_like(d)
                self.comm.reduce_scatter_into(out, dn)
                if self.residual:
                    out = out + d
                dn = out
            if grads is not None:
                self._update(i, grads)
            d = dn
        full_y = torch.empty((T, xs.shape[1]), dtype=xs.dtype, device=xs.device)
        self.comm.all_gather_into(full_y, xs)
        return full_y

    def _update(self, layer: int, grads) -> None:
        for j, (p, g) in enumerate(zip(self.shards[layer], grads)):
            self.opt.update(layer * 4 + j, p, g)

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        self.opt.begin_step()
        return self._step_sp(x, dloss_dx) if self.sp else self._step_ar(x, dloss_dx)

    def run(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> None:
        for x, dy in batches:
            self.step(x, dy)

    def full_params(self) -> List[Params]:
        """Reassemble: W1/b1 along dim 0, W2 along dim 1 (``train_ffns.py:335-338``); b2 is replicated."""
        from .base import tp_shard_dim
        out = []
        for lp in self.shards:
            fl = []
            for j, s in enumerate(lp):
                dim = tp_shard_dim(j)
                if dim is None or self.world == 1:
                    fl.append(s)
                    continue
                src = s if dim == 0 else s.t().contiguous()
                full = torch.empty((src.shape[0] * self.world, *src.shape[1:]), dtype=s.dtype, device=s.device)
                self.comm.all_gather_into(full, src.contiguous())
                fl.append(full if dim == 0 else full.t().contiguous())
            out.append(fl)
        return out
