"""Op interface on the hand-written sm_100a kernels (``libb200ffn.so``).

bf16 tensors whose shapes meet the TMA constraints go to the tcgen05 GEMM with the elementwise work fused in
its epilogue.  fp32 tensors (the reference's dtype and the CLI default) go to the SAME mainloop through split operands:
three bf16 planes per matrix, six plane products accumulated in fp32 (``csrc/gemm_f32.cu``), fp32 epilogues.  Anything
else (toy widths such as the CLI default ``model_size=4``) goes to the ``TorchOps`` path and is counted in ``fallbacks``
so benchmarks can assert it stayed at zero."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import native as nv
from .torch_ops import TorchOps


class CudaOps(TorchOps):
    name = "native"
    fused_update = True

    def __init__(self, max_ctas: int = 0) -> None:
        nv.load()
        self.max_ctas = max_ctas
        self.fallbacks = 0
        self.relu_mask = True

    # ---- helpers -----------------------------------------------------------------------------------------
    @staticmethod
    def _ok(*tensors: torch.Tensor) -> bool:
        for t in tensors:
            if t is None:
                continue
            if (not t.is_cuda or t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1
                    or t.stride(0) % 8 or t.data_ptr() % 16):
                return False
        return True

    @staticmethod
    def _bias32(bias: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """The epilogues take an fp32 bias vector; biases are stored fp32, so this is normally the tensor itself."""
        if bias is None or (bias.dtype == torch.float32 and bias.is_contiguous()):
            return bias
        return bias.float().contiguous()

    def _fb(self):
        self.fallbacks += 1
        return super()

    # ---- fp32 (reference-parity) path: split operands, six bf16 tensor-core products, fp32 epilogues ------------------
    @staticmethod
    def _ok32(*tensors: torch.Tensor) -> bool:
        for t in tensors:
            if t is None:
                continue
            if (not t.is_cuda or t.dtype != torch.float32 or t.dim() != 2 or t.stride(1) != 1
                    or t.stride(0) % 8 or t.shape[1] % 8 or t.data_ptr() % 16):
                return False
        return True

    @staticmethod
    def _is32(*tensors: torch.Tensor) -> bool:
        return all(t is None or t.dtype == torch.float32 for t in tensors)

    # ---- forward -----------------------------------------------------------------------------------------
    def linear_act(self, x, W, act, bias=None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        T, K = x.shape
        N = W.shape[0]
        if self._is32(x, W) and self._ok32(x, W):
            a = torch.empty(T, N, dtype=torch.float32, device=x.device)
            pre = torch.empty_like(a) if act == "gelu" else None
            nv.gemm_f32(nv.split3(x), nv.MAJOR_K, nv.split3(W), nv.MAJOR_K, a, T, N, K, nv.EPI_RELU if act == "relu" else nv.EPI_GELU,
                        out2=pre, bias=self._bias32(bias), max_ctas=self.max_ctas)
            return a, pre
        if not self._ok(x, W) or N % 8:
            return self._fb().linear_act(x, W, act, bias)
        a = torch.empty(T, N, dtype=torch.bfloat16, device=x.device)
        b32 = self._bias32(bias)
        if act == "relu":
            if self.relu_mask and N % 128 == 0:
                # also emit the packed sign mask (1 bit/element): the backward reads 1/16 of the bytes of `a`
                mask = nv.relu_mask_empty(T, N, x.device)
                nv.gemm(x, nv.MAJOR_K, W, nv.MAJOR_K, a, T, N, K, nv.EPI_RELU_MASK, out2=mask, bias=b32, max_ctas=self.max_ctas)
                return a, mask
            nv.gemm(x, nv.MAJOR_K, W, nv.MAJOR_K, a, T, N, K, nv.EPI_RELU, bias=b32, max_ctas=self.max_ctas)
            return a, None
        pre = torch.empty_like(a)
        nv.gemm(x, nv.MAJOR_K, W, nv.MAJOR_K, a, T, N, K, nv.EPI_GELU, out2=pre, bias=b32, max_ctas=self.max_ctas)
        return a, pre

    def linear(self, a, W, bias=None, residual=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        T, K = a.shape
        N = W.shape[0]
        if self._is32(a, W, residual) and self._ok32(a, W, residual):
            y = torch.empty(T, N, dtype=torch.float32, device=a.device) if out is None else out
            nv.gemm_f32(nv.split3(a), nv.MAJOR_K, nv.split3(W), nv.MAJOR_K, y, T, N, K, nv.EPI_ADD if residual is not None else nv.EPI_STORE,
                        aux=residual, bias=self._bias32(bias), max_ctas=self.max_ctas)
            return y
        if not self._ok(a, W, residual) or N % 8:
            return self._fb().linear(a, W, bias, residual)
        y = torch.empty(T, N, dtype=torch.bfloat16, device=a.device) if out is None else out
        b32 = self._bias32(bias)
        if residual is not None:
            nv.gemm(a, nv.MAJOR_K, W, nv.MAJOR_K, y, T, N, K, nv.EPI_ADD, aux=residual, bias=b32, max_ctas=self.max_ctas)
        else:
            nv.gemm(a, nv.MAJOR_K, W, nv.MAJOR_K, y, T, N, K, nv.EPI_STORE, bias=b32, max_ctas=self.max_ctas)
        return y

    # ---- backward ----------------------------------------------------------------------------------------
    def grad_input(self, dy, W, act=None, aux=None, add=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        T, K = dy.shape           # K = out features of the layer
        N = W.shape[1]            # in features
        mask_aux = aux is not None and aux.dtype == torch.uint8
        if self._is32(dy, W, aux, add) and self._ok32(dy, W, aux, add) and not (act is not None and add is not None):
            dx = torch.empty(T, N, dtype=torch.float32, device=dy.device) if out is None else out
            epi, ax = nv.EPI_STORE, None
            if act == "relu":
                epi, ax = nv.EPI_RELU_BWD, aux
            elif act == "gelu":
                epi, ax = nv.EPI_GELU_BWD, aux
            elif add is not None:
                epi, ax = nv.EPI_ADD, add
            nv.gemm_f32(nv.split3(dy), nv.MAJOR_K, nv.split3(W), nv.MAJOR_MN, dx, T, N, K, epi, aux=ax, max_ctas=self.max_ctas)
            return dx
        if not self._ok(dy, W, None if mask_aux else aux, add) or N % 8 or (act is not None and add is not None):
            if mask_aux:
                raise nv.NativeError("grad_input: a packed ReLU mask can only be consumed by the native kernel")
            return self._fb().grad_input(dy, W, act, aux, add)
        dx = torch.empty(T, N, dtype=torch.bfloat16, device=dy.device) if out is None else out
        if act == "relu" and aux.dtype == torch.uint8:
            nv.gemm(dy, nv.MAJOR_K, W, nv.MAJOR_MN, dx, T, N, K, nv.EPI_RELU_BWD_MASK, aux=aux, max_ctas=self.max_ctas)
        elif act == "relu":
            nv.gemm(dy, nv.MAJOR_K, W, nv.MAJOR_MN, dx, T, N, K, nv.EPI_RELU_BWD, aux=aux, max_ctas=self.max_ctas)
        elif act == "gelu":
            nv.gemm(dy, nv.MAJOR_K, W, nv.MAJOR_MN, dx, T, N, K, nv.EPI_GELU_BWD, aux=aux, max_ctas=self.max_ctas)
        elif add is not None:
            nv.gemm(dy, nv.MAJOR_K, W, nv.MAJOR_MN, dx, T, N, K, nv.EPI_ADD, aux=add, max_ctas=self.max_ctas)
        else:
            nv.gemm(dy, nv.MAJOR_K, W, nv.MAJOR_MN, dx, T, N, K, nv.EPI_STORE, max_ctas=self.max_ctas)
        return dx

    def grad_weight(self, dy, x, out_dtype=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        T, M = dy.shape
        N = x.shape[1]
        if self._is32(dy, x) and self._ok32(dy, x) and T % 64 == 0 and out_dtype in (None, torch.float32):
            dW = out if out is not None else torch.empty(M, N, dtype=torch.float32, device=dy.device)
            nv.gemm_f32(nv.split3(dy), nv.MAJOR_MN, nv.split3(x), nv.MAJOR_MN, dW, M, N, T, nv.EPI_STORE, max_ctas=self.max_ctas)
            return dW
        if not self._ok(dy, x) or N % 8 or M % 8:
            return self._fb().grad_weight(dy, x, out_dtype)
        f32 = out_dtype == torch.float32
        dW = out if out is not None else torch.empty(M, N, dtype=torch.float32 if f32 else torch.bfloat16, device=dy.device)
        nv.gemm(dy, nv.MAJOR_MN, x, nv.MAJOR_MN, dW, M, N, T, nv.EPI_STORE_F32 if f32 else nv.EPI_STORE,
                max_ctas=self.max_ctas)
        return dW

    def grad_weight_sgd_(self, W, dy, x, lr) -> None:
        """``W -= lr * dy^T @ x`` in ONE kernel: the gradient tile never leaves TMEM/registers."""
        T, M = dy.shape
        N = x.shape[1]
        if self._is32(dy, x, W) and self._ok32(dy, x, W) and T % 64 == 0:
            nv.gemm_f32(nv.split3(dy), nv.MAJOR_MN, nv.split3(x), nv.MAJOR_MN, W, M, N, T, nv.EPI_SGD, alpha=lr, max_ctas=self.max_ctas)
            return
        if not self._ok(dy, x, W) or N % 8 or M % 8:
            return self._fb().grad_weight_sgd_(W, dy, x, lr)
        nv.gemm(dy, nv.MAJOR_MN, x, nv.MAJOR_MN, W, M, N, T, nv.EPI_SGD, alpha=lr, max_ctas=self.max_ctas)

    def grad_bias(self, dy: torch.Tensor) -> torch.Tensor:
        """``db = dy.sum(0)`` in fp32: the native two-stage column sum (reads ``dy`` once)."""
        if self._ok(dy) and dy.shape[1] % 8 == 0:
            return nv.colsum(dy)
        return self._fb().grad_bias(dy)

    # ---- optimizer ---------------------------------------------------------------------------------------
    def sgd_(self, p, g, lr) -> None:
        if p.is_cuda and p.dtype == torch.float32 and g.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous():
            nv.sgd_f32_(p, g, lr)           # fp32 parameters: the biases
            return
        if p.is_cuda and p.dtype == torch.bfloat16 and p.is_contiguous() and g.is_contiguous() \
                and g.dtype in (torch.bfloat16, torch.float32) and p.data_ptr() % 16 == 0 and g.data_ptr() % 16 == 0:
            nv.sgd_(p, g, lr)
        else:
            self._fb().sgd_(p, g, lr)

    def adam_(self, p, g, m, v, lr, beta1, beta2, eps, step) -> None:
        if p.is_cuda and p.dtype == torch.float32 and g.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous():
            nv.adam_f32_(p, g, m, v, lr, beta1, beta2, eps, step)
            return
        if p.is_cuda and p.dtype == torch.bfloat16 and p.is_contiguous() and g.is_contiguous() \
                and g.dtype in (torch.bfloat16, torch.float32):
            nv.adam_(p, g, m, v, lr, beta1, beta2, eps, step)
        else:
            self._fb().adam_(p, g, m, v, lr, beta1, beta2, eps, step)

    def sumsq(self, x):
        if x.is_cuda and x.dtype == torch.bfloat16 and x.is_contiguous() and x.data_ptr() % 16 == 0:
            return nv.sumsq(x)
        return self._fb().sumsq(x)
