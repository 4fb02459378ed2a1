"""The model: a stack of Transformer FFN sub-blocks with a hand-written forward/backward schedule.

Reference semantics (``train_ffns.py``): parameters are ``[W1 (F x D), W2 (D x F)]`` per layer in the
``nn.Linear`` ``[out, in]`` layout (``:35-39``); forward is ``relu(x W1^T) W2^T`` (``:54-58``); backward
recomputes the hidden activation from the saved layer input (``:61-70``); the stack saves only layer inputs
(``:72-81``) and walks the layers in reverse calling a per-layer communication hook (``:83-94``).

Extensions: optional bias, GELU, residual, and a *keep-activations* policy (default) that
stores ``a = act(h)`` in forward instead of recomputing it - 16 GiB for the 32-block bf16 benchmark stack, which a
180 GB part holds easily, and it removes one of the seven GEMMs per layer-step.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence, Tuple

import torch
from torch import nn

from ..config import INIT_SCALE
from ..utils.profiler import block_range

Params = List[torch.Tensor]            # [W1, W2] or [W1, W2, b1, b2]
LayerGrads = Tuple[torch.Tensor, ...]


# ------------------------------------------------------------------------------------------- init
def init_linear_layer(m: int, n: int, gen: Optional[torch.Generator], scale: float = INIT_SCALE) -> torch.Tensor:
    """``scale * randn(n, m)``: a weight for ``m -> n`` features, ``[out, in]`` (``train_ffns.py:35-36``)."""
    return scale * torch.randn((n, m), generator=gen)


def init_tlayer_ffn(emb_dim: int, ffn_dim: int, gen: Optional[torch.Generator], bias: bool = False) -> Params:
    p = [init_linear_layer(emb_dim, ffn_dim, gen), init_linear_layer(ffn_dim, emb_dim, gen)]
    if bias:
        p += [torch.zeros(ffn_dim), torch.zeros(emb_dim)]          # fp32, and kept fp32 on every path
    return p


def init_stack(layers: int, emb_dim: int, ffn_dim: int, gen: Optional[torch.Generator], bias: bool = False
               ) -> List[Params]:
    return [init_tlayer_ffn(emb_dim, ffn_dim, gen, bias) for _ in range(layers)]


def num_params(layers_params: Sequence[Params]) -> int:
    return sum(p.numel() for lp in layers_params for p in lp)


# ------------------------------------------------------------------------------------------- state_dict layout
class FFNBlock(nn.Module):
    def __init__(self, emb_dim: int, ffn_dim: int, bias: bool = False):
        super().__init__()
        self.ffn1 = nn.Linear(emb_dim, ffn_dim, bias=bias)
        self.ffn2 = nn.Linear(ffn_dim, emb_dim, bias=bias)


class FFNStack(nn.Module):
    """Plain ``nn.Module`` whose ``state_dict`` is the checkpoint layout: ``layers.{i}.ffn1.weight`` ``[F, D]``,
    ``layers.{i}.ffn2.weight`` ``[D, F]`` (+ ``.bias``).  Its autograd forward is the oracle for the manual
    backward."""

    def __init__(self, layers: int, emb_dim: int, ffn_dim: int, bias: bool = False, activation: str = "relu",
                 residual: bool = False):
        super().__init__()
        self.layers = nn.ModuleList([FFNBlock(emb_dim, ffn_dim, bias) for _ in range(layers)])
        self.activation, self.residual = activation, residual

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        act = torch.relu if self.activation == "relu" else torch.nn.functional.gelu
        for blk in self.layers:
            y = blk.ffn2(act(blk.ffn1(x)))
            x = x + y if self.residual else y
        return x

    @staticmethod
    def from_params(layers_params: Sequence[Params], activation: str = "relu", residual: bool = False) -> "FFNStack":
        F_, D = layers_params[0][0].shape
        m = FFNStack(len(layers_params), D, F_, bias=len(layers_params[0]) == 4, activation=activation, residual=residual)
        m.load_state_dict(params_to_state_dict(layers_params))
        return m


def params_to_state_dict(layers_params: Sequence[Params]) -> dict:
    sd = {}
    for i, lp in enumerate(layers_params):
        sd[f"layers.{i}.ffn1.weight"] = lp[0].detach()
        sd[f"layers.{i}.ffn2.weight"] = lp[1].detach()
        if len(lp) == 4:
            sd[f"layers.{i}.ffn1.bias"] = lp[2].detach()
            sd[f"layers.{i}.ffn2.bias"] = lp[3].detach()
    return sd


def state_dict_to_params(sd: dict) -> List[Params]:
    n = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("layers."))
    out = []
    for i in range(n):
        lp = [sd[f"layers.{i}.ffn1.weight"], sd[f"layers.{i}.ffn2.weight"]]
        if f"layers.{i}.ffn1.bias" in sd:
            lp += [sd[f"layers.{i}.ffn1.bias"], sd[f"layers.{i}.ffn2.bias"]]
        out.append(lp)
    return out


# ------------------------------------------------------------------------------------------- block fwd / bwd
@dataclass
class BlockOpts:
    activation: str = "relu"
    residual: bool = False
    recompute: bool = False     # True = reference behaviour (recompute h in backward, train_ffns.py:63)


@dataclass
class Saved:
    """What one block keeps for backward: always its input (the reference's ``acts``), optionally a / pre-act."""
    x: torch.Tensor
    a: Optional[torch.Tensor] = None
    pre: Optional[torch.Tensor] = None


def _bias(p: Params, j: int) -> Optional[torch.Tensor]:
    return p[2 + j] if len(p) == 4 else None


def tlayer_ffn_fwd(ops, p: Params, x: torch.Tensor, opts: BlockOpts = BlockOpts()) -> Tuple[torch.Tensor, Saved]:
    """``y = act(x W1^T + b1) W2^T + b2 (+ x)`` (``train_ffns.py:54-58``)."""
    with block_range("ffn_fwd"):
        a, pre = ops.linear_act(x, p[0], opts.activation, _bias(p, 0))
        y = ops.linear(a, p[1], _bias(p, 1), residual=x if opts.residual else None)
    return y, (Saved(x) if opts.recompute else Saved(x, a, pre))


def tlayer_ffn_bkwd(ops, dy: torch.Tensor, p: Params, saved: Saved, opts: BlockOpts = BlockOpts(),
                    need_dx: bool = True, update_lr: Optional[float] = None
                    ) -> Tuple[Optional[torch.Tensor], Optional[LayerGrads]]:
    """Backward of one block (``train_ffns.py:61-70``).  Returns ``(dx, (dW1, dW2[, db1, db2]))``.

    The input-gradient GEMM of each linear is issued BEFORE its weight-gradient GEMM so that, when
    ``update_lr`` is given, the weight can be updated in place inside the dW GEMM's epilogue (``W -= lr dW``,
    one kernel, no gradient tensor) - the returned weight grads are then ``None`` (with bias: ``(None, None, db1, db2)``,
    without: ``None``).  Biases are fp32 whatever the weights' dtype, and so are their gradients.  ``need_dx=False`` skips the
    layer-0 input gradient the reference computes and discards (``:88``).
    """
    with block_range("ffn_bwd"):
        return _tlayer_ffn_bkwd(ops, dy, p, saved, opts, need_dx, update_lr)


def _tlayer_ffn_bkwd(ops, dy, p, saved, opts, need_dx, update_lr):
    x = saved.x
    a, pre = saved.a, saved.pre
    if a is None:                                   # recompute policy (:63, :66)
        a, pre = ops.linear_act(x, p[0], opts.activation, _bias(p, 0))
    # ReLU backward needs only the sign of h: the native ops hand back a packed bitmask in `pre`, else use `a` itself
    aux = (pre if pre is not None else a) if opts.activation == "relu" else pre
    dh = ops.grad_input(dy, p[1], act=opts.activation, aux=aux)            # da then activation backward (:66-67)
    has_bias = len(p) == 4
    fused = update_lr is not None            # weights updated in the dW GEMM's epilogue; bias gradients are still returned
    if fused:
        ops.grad_weight_sgd_(p[1], dy, a, update_lr)
        dW2 = None
    else:
        dW2 = ops.grad_weight(dy, a)                                       # (:66)
    dx = None
    if need_dx:
        dx = ops.grad_input(dh, p[0], add=dy if opts.residual else None)   # (:68)
    if fused:
        ops.grad_weight_sgd_(p[0], dh, x, update_lr)
        dW1 = None
    else:
        dW1 = ops.grad_weight(dh, x)                                       # (:68)
    if has_bias:
        return dx, (dW1, dW2, ops.grad_bias(dh), ops.grad_bias(dy))
    return dx, (None if fused else (dW1, dW2))


# ------------------------------------------------------------------------------------------- stack fwd / bwd
def tlayers_ffn_fwd(ops, layers_params: Sequence[Params], x: torch.Tensor, opts: BlockOpts = BlockOpts(),
                    before_comms_hook: Optional[Callable] = None) -> Tuple[torch.Tensor, List[Saved]]:
    """Forward through the stack, saving each block's input (``train_ffns.py:72-81``).  ``before_comms_hook(i)``
    runs before block ``i`` (the reference's prefetch extension point, ``:78-79``)."""
    y = x
    acts: List[Saved] = []
    for i, p in enumerate(layers_params):
        if before_comms_hook is not None:
            p = before_comms_hook(i) or p
        y, s = tlayer_ffn_fwd(ops, p, y, opts)
        acts.append(s)
    return y, acts


def tlayers_ffn_bkwd(ops, dy: torch.Tensor, layers_params: Sequence[Params], acts: Sequence[Saved],
                     opts: BlockOpts = BlockOpts(), after_comms_hook: Optional[Callable] = None,
                     skip_first_dx: bool = True) -> Tuple[List[LayerGrads], List]:
    """Reverse walk (``train_ffns.py:83-94``): per layer, compute grads then call ``after_comms_hook(i, grads)``
    (the DDP per-layer bucket trigger, ``:90-91``).  Returns grads and hook results in layer order."""
    grads_lst: List = []
    handles: List = []
    d = dy
    for i in reversed(range(len(layers_params))):
        d, g = tlayer_ffn_bkwd(ops, d, layers_params[i], acts[i], opts, need_dx=not (skip_first_dx and i == 0))
        grads_lst.append(g)
        handles.append(after_comms_hook(i, g) if after_comms_hook is not None else None)
    return list(reversed(grads_lst)), list(reversed(handles))
