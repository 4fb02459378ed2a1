"""Pure-PyTorch implementation of the op interface: the CPU/gloo plumbing path, the numerical oracle for the
native kernels, and the cuBLAS/ATen baseline on GPU.  Semantics follow the reference's L1 ops:
``linear_fwd`` (train_ffns.py:41-42), ``t_linear_bkwd`` (:44-45), ``t_relu_fwd`` (:47-48), ``t_relu_bkwd_``
(:50-52) and the SGD expressions (:114, :172, :259, :312)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F


class TorchOps:
    name = "torch"
    fused_update = False    # grad_weight_sgd_ is a matmul + add_, not one kernel

    # ---- forward -------------------------------------------------------------------------------------
    def linear_act(self, x: torch.Tensor, W: torch.Tensor, act: str, bias: Optional[torch.Tensor] = None
                   ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """``a = act(x @ W^T + b)``; returns ``(a, pre)`` where ``pre`` is the pre-activation when the backward
        needs it (GELU) and ``None`` for ReLU (its mask is recoverable from ``a > 0``)."""
        h = torch.matmul(x, W.t())
        if bias is not None:
            h = h + bias.to(h.dtype)
        if act == "relu":
            return torch.where(torch.le(h, 0), torch.zeros((), dtype=h.dtype, device=h.device), h), None
        return F.gelu(h), h

    def linear(self, a: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor] = None,
               residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        y = torch.matmul(a, W.t())
        if bias is not None:
            y = y + bias.to(y.dtype)
        if residual is not None:
            y = y + residual
        return y

    # ---- backward ------------------------------------------------------------------------------------
    def grad_input(self, dy: torch.Tensor, W: torch.Tensor, act: Optional[str] = None,
                   aux: Optional[torch.Tensor] = None, add: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``dx = dy @ W`` then, optionally, the activation backward (``aux`` = ReLU output or GELU
        pre-activation) or a residual-gradient add."""
        dx = torch.matmul(dy, W)
        if act == "relu":
            dx = dx.masked_fill_(aux <= 0, 0)
        elif act == "gelu":
            ct = torch.float64 if aux.dtype == torch.float64 else torch.float32
            a32 = aux.to(ct)
            cdf = 0.5 * (1.0 + torch.erf(a32 * 0.7071067811865476))
            pdf = torch.exp(-0.5 * a32 * a32) * 0.3989422804014327
            dx = (dx.to(ct) * (cdf + a32 * pdf)).to(dy.dtype)
        if add is not None:
            dx = dx + add
        return dx

    def grad_weight(self, dy: torch.Tensor, x: torch.Tensor, out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
        """``dW[out,in] = dy^T @ x`` (einsum 'bc,bd->cd', train_ffns.py:45)."""
        dW = torch.matmul(dy.t(), x)
        return dW if out_dtype is None else dW.to(out_dtype)

    def grad_weight_sgd_(self, W: torch.Tensor, dy: torch.Tensor, x: torch.Tensor, lr: float) -> None:
        W.add_(torch.matmul(dy.t(), x), alpha=-lr)

    def grad_bias(self, dy: torch.Tensor) -> torch.Tensor:
        return dy.float().sum(0)

    # ---- optimizer -----------------------------------------------------------------------------------
    def sgd_(self, p: torch.Tensor, g: torch.Tensor, lr: float) -> None:
        p.add_(g.to(p.dtype), alpha=-lr)

    def adam_(self, p, g, m, v, lr, beta1, beta2, eps, step) -> None:
        g32 = g.float()
        m.mul_(beta1).add_(g32, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g32, g32, value=1 - beta2)
        bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
        upd = (m / bc1) / ((v / bc2).sqrt() + eps)
        p.copy_((p.float() - lr * upd).to(p.dtype))

    def sumsq(self, x: torch.Tensor) -> torch.Tensor:
        return (x.float() ** 2).sum().reshape(1)
