"""FSDP with the forward parameter all-gather INSIDE the GEMM that consumes the gathered weight.

The all-gather + GEMM kernel (``csrc/gather_gemm.cu``) gathers a row-sharded, K-major **A** operand.  FSDP shards every
weight along dim 0 (``train_ffns.py:265-266``), i.e. along its rows, so the weight can be that A operand if each
forward GEMM is written as ``W @ act^T``: activations are kept TRANSPOSED, ``[features, tokens]``, for the whole step.

    forward   a^T [F,T] = act(W1 [F,D] @ x^T [D,T])        W1 gathered in the kernel      (:55-56 + :203-209)
              y^T [D,T] = W2 [D,F] @ a^T [F,T] (+ x^T)     W2 gathered in the kernel      (:57)
    backward  dh^T [F,T] = act'(W2^T @ dy^T)                W2 gathered in the kernel      (:66-67 + :244-249)
              W2 shard  -= lr * rows of (dy^T @ a)          fused reduce-scatter + SGD/Adam (:253-259)
              dx^T [D,T] = W1^T @ dh^T (+ dy^T)             W1 gathered in the kernel      (:68)
              W1 shard  -= lr * rows of (dh^T @ x)

Every product is one of the operand-layout combinations the tcgen05 kernel already runs (K- or MN-major A and B).  In
the backward pass the weight is read MN-major and its sharded dimension is the CONTRACTED one: the kernel's second form
(K-sharded A) starts every tile's k-loop at the rank's own shard and waits per k-block for the rest.  The last two
layers' weights are still complete from the forward pass and are used as they are (as the reference reuses the last
layer's, ``:244``).  No copy-engine prefetch is left in this trainer.  The schedule is ``fsdp_t_step`` below, written
against three primitives so that it is checked on CPU (``tests/test_fsdp_t_schedule_cpu.py``).

Validated on B200 in round 2 (2 GPUs, ``profiles/tests/r2_call1_2gpu``); selected with ``TrainConfig.fsdp_gather = "fused"`` (``--fsdp_gather fused``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..ops import native as nv
from .base import set_update_rounding
from .fused import FOP_SGD, OWN_ROW_SHARDED, FusedFSDPTrainer, GatherChannel

K_, MN_ = nv.MAJOR_K, nv.MAJOR_MN


def fsdp_t_step(t, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
    """One optimizer step in transposed-activation space.  ``t`` provides ``L, opts`` and

    * ``t.gather_w_gemm(l, k, B, N, epi, aux)``   ``epi(W[l][k] @ B)`` with ``W``'s row shards gathered in the kernel;
      ``B`` is ``[K, N]`` (MN-major); returns ``(out [rows(W), N], out2)`` and leaves the full ``W`` for the backward pass
    * ``t.bwd_w_gemm(l, k, B, N, epi, aux)``   ``epi(W[l][k]^T @ B)``: ``W`` read MN-major, sharded along the contracted
      dimension; gathered in the kernel unless it is still complete from the forward pass; returns ``out [cols(W), N]``
    * ``t.transpose(x)``   contiguous ``x.T`` (native tiled transpose on the GPU)
    * ``t.update(l, k, A, B)``   ``W[l][k] shard -= lr * (rows of A @ B^T summed over ranks)``, A ``[M, T]``, B ``[N, T]``
    """
    act, residual, L = t.opts.activation, t.opts.residual, t.L
    T = x.shape[0]
    xT = t.transpose(x)                                                   # [D, T]
    saved = []
    for l in range(L):
        fwd = {"relu": nv.EPI_RELU_MASK if T % 128 == 0 else nv.EPI_RELU, "gelu": nv.EPI_GELU}[act]
        aT, aux = t.gather_w_gemm(l, 0, xT, T, fwd, None)                 # [F, T]
        if aux is None:
            aux = aT                                                      # plain ReLU: the output itself carries the sign
        yT, _ = t.gather_w_gemm(l, 1, aT, T, nv.EPI_ADD if residual else nv.EPI_STORE, xT if residual else None)   # [D, T]
        saved.append((xT, aT, aux))
        xT = yT
    y = t.transpose(xT)
    dT = t.transpose(dloss_dx)                                            # [D, T]
    for i in reversed(range(L)):
        xin, aT, aux = saved[i]
        bwd = (nv.EPI_RELU_BWD_MASK if aux.dtype == torch.uint8 else nv.EPI_RELU_BWD) if act == "relu" else nv.EPI_GELU_BWD
        # dh^T[f,t] = act'( sum_d W2[d,f] dy^T[d,t] ): A = W2 read MN-major (storage [K=D, M=F]), B = dy^T MN-major ([K=D, N=T])
        dhT = t.bwd_w_gemm(i, 1, dT, T, bwd, aux)
        t.update(i, 1, dT, aT)                                            # dW2[d,f] = sum_t dy^T[d,t] a^T[f,t]
        dn = None
        if i > 0:
            # dx^T[d,t] = sum_f W1[f,d] dh^T[f,t] (+ dy^T): A = W1 read MN-major ([K=F, M=D]), B = dh^T MN-major ([K=F, N=T])
            dn = t.bwd_w_gemm(i, 0, dhT, T, nv.EPI_ADD if residual else nv.EPI_STORE, dT if residual else None)
        t.update(i, 0, dhT, xin)                                          # dW1[f,d] = sum_t dh^T[f,t] x^T[d,t]
        saved[i] = None
        dT = dn
    return y


class FusedFSDPGatherTrainer(FusedFSDPTrainer):
    """``FusedFSDPTrainer``'s fused reduce-scatter + sharded optimizer step, with every parameter all-gather inside the
    GEMM that consumes the gathered weight (forward: row-sharded A; backward: K-sharded A)."""

    comm_name = "fused: parameter all-gather inside the consuming GEMMs (forward and backward); dW GEMM + reduce-scatter + sharded update in one kernel"

    def __init__(self, ops, comm, cfg, shards, device):
        super().__init__(ops, comm, cfg, shards, device)
        if self.adam:
            self.comm_name = self.comm_name.replace("sharded update", "sharded Adam")
        self.device = device
        self.gchan = GatherChannel(self.sw, max(cfg.model_size, cfg.ffn_dim))
        self.sw.host_barrier()

    # ---- primitives of fsdp_t_step ---------------------------------------------------------------------------
    def _peers_of_weight(self, l: int, k: int, full: torch.Tensor):
        """Per-rank base pointers such that ``base[q] + q * shard_bytes`` is rank q's shard: peers' point into their shard
        storage (never dereferenced outside their own rows), mine is the full buffer the kernel completes."""
        sh = self.shards[l][k]
        sb = sh.numel() * 2
        ptrs = [full.data_ptr() if q == self.rank else self.sbuf.peer_ptrs[q] + self.shard_off[l][k] - q * sb for q in range(self.world)]
        return (C.c_void_p * self.world)(*ptrs)

    def gather_w_gemm(self, l: int, k: int, B: torch.Tensor, N: int, epi: int, aux: Optional[torch.Tensor]):
        full = self.full[l % 2][k]
        sh = self.shards[l][k]
        rows = sh.shape[0]
        full[self.rank * rows:(self.rank + 1) * rows].copy_(sh)           # my rows; the kernel pulls the others
        M, K = full.shape
        out = torch.empty(M, N, dtype=torch.bfloat16, device=self.device)
        out2 = None
        if epi == nv.EPI_RELU_MASK:
            out2 = nv.relu_mask_empty(M, N, self.device)
        elif epi == nv.EPI_GELU:
            out2 = torch.empty_like(out)
        self.gchan.gemm(self._peers_of_weight(l, k, full), B, MN_, out, M, N, K, epi, aux=aux, out2=out2)
        return out, out2

    def bwd_w_gemm(self, l: int, k: int, B: torch.Tensor, N: int, epi: int, aux: Optional[torch.Tensor]):
        full = self.full[l % 2][k]
        Kd, M = full.shape                                                # read MN-major: storage [K, M]
        out = torch.empty(M, N, dtype=torch.bfloat16, device=self.device)
        if l >= self.L - 2:                                               # still complete from this step's forward kernels
            return nv.gemm(full, MN_, B, MN_, out, M, N, Kd, epi, aux=aux)
        sh = self.shards[l][k]
        rows = sh.shape[0]
        full[self.rank * rows:(self.rank + 1) * rows].copy_(sh)
        self.gchan.gemm(self._peers_of_weight(l, k, full), B, MN_, out, M, N, Kd, epi, aux=aux, a_major=MN_)
        return out

    def transpose(self, x: torch.Tensor) -> torch.Tensor:
        return nv.transpose(x) if (x.is_cuda and x.dtype == torch.bfloat16 and x.stride(1) == 1 and x.stride(0) % 2 == 0 and x.shape[0] % 2 == 0) else x.t().contiguous()

    def update(self, l: int, k: int, A: torch.Tensor, B: torch.Tensor) -> None:
        M, T = A.shape
        N = B.shape[0]
        out, row0 = self.local_out[2 * l + k], self.rank * (M // self.world)
        if self.adam:
            self.chan.gemm_adam(A, K_, B, K_, M, N, T, out, N, row0, OWN_ROW_SHARDED, False, self.state[l][k], self.lr, self.t)
        else:
            self.chan.gemm(A, K_, B, K_, M, N, T, out, N, row0, OWN_ROW_SHARDED, FOP_SGD, False, self.lr)

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        self.t += 1
        set_update_rounding(self.cfg, self.ops, self.t)
        return fsdp_t_step(self, x, dloss_dx)
