"""ctypes binding of the in-tree native library (``_native/libb200ffn.so``, C ABI in ``include/b200ffn.h``).

The wrappers take torch CUDA tensors, pass raw device pointers plus the *current* torch stream, and raise
``NativeError`` with the library's message on failure - there is no silent fallback: on a GPU box a missing
or broken library is an error, not a reason to run eager PyTorch.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path
from typing import Optional

import torch

import os

# B200_NATIVE_LIB selects an A/B build of the same sources (build_native.build_variant); default = the in-tree library
_LIB_PATH = Path(os.environ.get("B200_NATIVE_LIB") or (Path(__file__).resolve().parent.parent / "_native" / "libb200ffn.so"))
_lib = None
_lock = threading.Lock()

MAJOR_K = 0
MAJOR_MN = 1

EPI_STORE, EPI_RELU, EPI_GELU, EPI_RELU_BWD, EPI_GELU_BWD, EPI_ADD, EPI_SGD, EPI_STORE_F32, EPI_RELU_MASK, EPI_RELU_BWD_MASK = range(10)

#: number of kernels this process launched through the native library (bench.py reports it)
launch_count = 0


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    """True when the shared object exists (it may still fail to run without a B200)."""
    return _LIB_PATH.exists()


def _declare(lib) -> None:
    vp, ll, i, f = C.c_void_p, C.c_longlong, C.c_int, C.c_float
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_version.restype = i
    lib.b200_num_sms.restype = i
    lib.b200_device_error.argtypes = [C.POINTER(C.c_uint)]
    lib.b200_gemm_bf16.argtypes = [vp, ll, i, vp, ll, i, vp, ll, i, i, i, i, vp, ll, vp, ll, vp, f, i, vp]
    lib.b200_sgd_bf16.argtypes = [vp, vp, ll, f, i, vp]
    lib.b200_adam_bf16.argtypes = [vp, vp, vp, vp, ll, f, f, f, f, i, i, vp]
    lib.b200_relu_bf16.argtypes = [vp, vp, ll, vp]
    lib.b200_relu_bwd_bf16.argtypes = [vp, vp, ll, vp]
    lib.b200_add_bf16.argtypes = [vp, vp, vp, ll, vp]
    lib.b200_randn_bf16.argtypes = [vp, ll, C.c_ulonglong, f, vp]
    lib.b200_cast_f32_to_bf16.argtypes = [vp, vp, ll, vp]
    lib.b200_cast_bf16_to_f32.argtypes = [vp, vp, ll, vp]
    lib.b200_sumsq_scratch_bytes.restype = ll
    lib.b200_sumsq_bf16.argtypes = [vp, vp, ll, vp, ll, vp]
    lib.b200_fill_zero.argtypes = [vp, ll, vp]


def load():
    """Load (building first if the .so is missing and nvcc is present) and return the ctypes handle."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not _LIB_PATH.exists():
                from ..build_native import build
                build()
            lib = C.CDLL(str(_LIB_PATH), mode=C.RTLD_GLOBAL)
            _declare(lib)
            _optional_declare(lib)
            _lib = lib
    return _lib


def _optional_declare(lib) -> None:
    """Symbols added by later translation units (collectives, fused kernels); declared when present."""
    vp, ll, i, f = C.c_void_p, C.c_longlong, C.c_int, C.c_float
    pp = C.POINTER(C.c_void_p)
    u = C.c_uint
    table = {
        "b200_ipc_handle_bytes": ([], i),
        "b200_symm_alloc": ([pp, ll], i),
        "b200_symm_free": ([vp], i),
        "b200_ipc_get_handle": ([vp, vp], i),
        "b200_ipc_open_handle": ([pp, vp], i),
        "b200_ipc_close_handle": ([vp], i),
        "b200_enable_peer_access": ([i], i),
        "b200_coll_flag_bytes": ([], ll),
        "b200_barrier": ([pp, i, i, u, vp], i),
        "b200_all_reduce_bf16": ([pp, ll, ll, pp, i, i, u, i, vp], i),
        "b200_all_gather": ([pp, ll, ll, pp, i, i, u, i, vp], i),
        "b200_all_reduce_f32": ([pp, ll, ll, pp, i, i, u, i, vp], i),
        "b200_reduce_scatter_f32": ([pp, ll, ll, pp, i, i, u, i, vp], i),
        "b200_reduce_scatter_bf16": ([pp, ll, ll, pp, i, i, u, i, vp], i),
        "b200_all_gather_into": ([vp, pp, ll, ll, pp, i, i, u, i, vp], i),
        "b200_sumsq_static_max_elems": ([], ll),
        "b200_sumsq_bf16_static": ([vp, vp, ll, vp], i),
        "b200_all_gather_into_dma": ([vp, pp, ll, ll, pp, i, i, u, vp], i),
        "b200_fd_listen": ([C.c_char_p], i),
        "b200_fd_send": ([C.c_char_p, i, i], i),
        "b200_fd_recv": ([i, C.POINTER(i)], i),
        "b200_fd_close": ([i], i),
        "b200_vmm_query": ([C.POINTER(i)], i),
        "b200_vmm_granularity": ([i, i], ll),
        "b200_vmm_alloc": ([ll, pp, C.POINTER(C.c_ulonglong), C.POINTER(i)], i),
        "b200_vmm_import": ([i, ll, pp, C.POINTER(C.c_ulonglong)], i),
        "b200_vmm_free": ([vp, ll, C.c_ulonglong], i),
        "b200_mc_create": ([i, ll, C.POINTER(C.c_ulonglong), C.POINTER(i)], i),
        "b200_mc_import": ([i, C.POINTER(C.c_ulonglong)], i),
        "b200_mc_attach": ([C.c_ulonglong, i, C.c_ulonglong, ll, pp], i),
        "b200_all_reduce_nvls_bf16": ([vp, ll, pp, i, i, u, i, vp], i),
        "b200_fused_inbox_bytes": ([i, i], ll),
        "b200_fused_flag_bytes": ([i, i], ll),
        "b200_fused_owned_tiles": ([i, i, i, i], i),
        "b200_gemm_allreduce_bf16": ([vp, ll, i, vp, ll, i, i, i, i, pp, ll, i, pp, pp, i, i, i, i, u, i, i, i, f, i, vp], i),
        "b200_fused_adam_state_bytes": ([i], ll),
        "b200_gemm_allreduce_adam_bf16": ([vp, ll, i, vp, ll, i, i, i, i, pp, ll, i, pp, pp, i, i, i, i, u, i, i,
                                          vp, vp, f, f, f, f, i, f, i, vp], i),
        "b200_gather_gemm_ready_bytes": ([i], ll),
        "b200_gather_gemm_bf16": ([pp, i, vp, ll, i, vp, ll, i, i, i, i, vp, ll, vp, ll, vp, f, pp, vp, i, i, u, i, vp], i),
        "b200_debug_gather_chunk": ([i, i, i, i, ll, C.POINTER(ll)], i),
        "b200_fused_config": ([i, i], i),
        "b200_transpose_bf16": ([vp, ll, vp, ll, i, i, vp], i),
        "b200_split_f32": ([vp, vp, ll, i, ll, vp], i),
        "b200_gemm_f32": ([vp, ll, i, vp, ll, i, vp, ll, i, i, i, i, vp, ll, vp, ll, vp, f, i, vp], i),
        "b200_set_update_rounding": ([u], i),
        "b200_fused_partial_bytes": ([i], ll),
        "b200_nvls_collective": ([i, i, vp, vp, ll, pp, i, i, u, i, vp], i),
        "b200_gemm_allreduce_nvls_bf16": ([vp, ll, i, vp, ll, i, i, i, i, pp, ll, i, vp, vp, vp, pp, i, i, i, i, u, i, i, i, f, i, vp], i),
        "b200_sgd_f32": ([vp, vp, ll, f, vp], i),
        "b200_adam_f32": ([vp, vp, vp, vp, ll, f, f, f, f, i, vp], i),
        "b200_colsum_scratch_bytes": ([i, i], ll),
        "b200_colsum_bf16": ([vp, vp, ll, i, i, i, vp, ll, vp], i),
        "b200_bias_rows_bf16": ([vp, ll, vp, ll, vp, i, i, vp], i),
        "b200_emu_table_bytes": ([], ll),
        "b200_emu_gemm_allreduce_bf16": ([i, pp, ll, i, pp, ll, i, i, i, i, pp, ll, pp, pp, i, i, u, i, i, i, f,
                                         pp, pp, f, f, f, f, i, f, i, vp, vp], i),
        "b200_emu_gather_gemm_bf16": ([i, pp, i, pp, ll, i, pp, ll, i, i, i, i, pp, ll, pp, ll, pp, pp, u, i, vp, vp], i),
    }
    for name, sig in table.items():
        if not hasattr(lib, name):
            continue
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = sig


def _check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed (rc={rc}): {msg}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def device_error() -> tuple[int, int, int, int]:
    buf = (C.c_uint * 4)()
    load().b200_device_error(buf)
    return tuple(int(x) for x in buf)  # type: ignore[return-value]


def num_sms() -> int:
    return int(load().b200_num_sms())


# --------------------------------------------------------------------------------------------- ReLU sign masks
# The RELU_MASK / RELU_BWD_MASK epilogues exchange 1 bit per element in a tile-blocked layout (csrc/gemm_kernel.cuh): the 16
# bytes of (row, 128-column chunk cc) sit at ((row // 128) * n_chunks + cc) * 2048 + (row % 128) * 16, so a warp's 32 rows
# are 512 contiguous bytes.  The buffer is a uint8 tensor of shape [ceil(M/128)*128, n_chunks*16] (the bytes of a row-major
# mask padded to whole tiles; only their ORDER differs).
def relu_mask_empty(M: int, N: int, device) -> torch.Tensor:
    return torch.empty((M + 127) // 128 * 128, (N + 127) // 128 * 16, dtype=torch.uint8, device=device)


def unpack_relu_mask(mask: torch.Tensor, M: int, N: int) -> torch.Tensor:
    """Blocked mask buffer -> bool ``[M, N]`` (tests, debugging)."""
    tm, ncc = (M + 127) // 128, (N + 127) // 128
    v = mask.reshape(-1)[: tm * ncc * 2048].view(tm, ncc, 128, 16)
    bits = (v.unsqueeze(-1) >> torch.arange(8, device=mask.device, dtype=torch.uint8)) & 1
    return bits.permute(0, 2, 1, 3, 4).reshape(tm * 128, ncc * 128)[:M, :N].bool()


def pack_relu_mask(keep: torch.Tensor) -> torch.Tensor:
    """bool ``[M, N]`` -> blocked mask buffer (tests)."""
    M, N = keep.shape
    tm, ncc = (M + 127) // 128, (N + 127) // 128
    full = torch.zeros(tm * 128, ncc * 128, dtype=torch.bool, device=keep.device)
    full[:M, :N] = keep
    w = (1 << torch.arange(8, device=keep.device, dtype=torch.int32))
    by = (full.view(tm, 128, ncc, 16, 8).int() * w).sum(-1).to(torch.uint8)          # [tm, 128, ncc, 16]
    return by.permute(0, 2, 1, 3).contiguous().view(tm * 128, ncc * 16)


# --------------------------------------------------------------------------------------------- GEMM
def gemm_supported(M: int, N: int, K: int, *lds: int) -> bool:
    """Shapes the tcgen05 kernel accepts (TMA pitch and the 16-byte epilogue vectors need multiples of 8)."""
    return M > 0 and N > 0 and K > 0 and N % 8 == 0 and all(ld % 8 == 0 for ld in lds)


def gemm(A: torch.Tensor, a_major: int, B: torch.Tensor, b_major: int, out: torch.Tensor, M: int, N: int, K: int,
         epilogue: int = EPI_STORE, aux: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None,
         bias: Optional[torch.Tensor] = None, alpha: float = 1.0, max_ctas: int = 0) -> torch.Tensor:
    """``out[M,N] = epilogue(sum_k A(m,k) B(n,k))``; see include/b200ffn.h for majors and epilogues."""
    global launch_count
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    assert A.stride(-1) == 1 and B.stride(-1) == 1 and out.stride(-1) == 1
    # (for the mask epilogues aux / out2 are uint8 [M, bytes]: stride(0) is then a pitch in bytes, as the C ABI wants)
    rc = load().b200_gemm_bf16(
        A.data_ptr(), A.stride(0), a_major, B.data_ptr(), B.stride(0), b_major, out.data_ptr(), out.stride(0),
        M, N, K, epilogue, _ptr(aux), aux.stride(0) if aux is not None else 0, _ptr(out2),
        out2.stride(0) if out2 is not None else 0, _ptr(bias), float(alpha), int(max_ctas), _stream())
    _check(rc, "b200_gemm_bf16")
    launch_count += 1
    return out


def split3(x: torch.Tensor) -> torch.Tensor:
    """fp32 ``[rows, cols]`` -> its three bf16 planes ``[3, rows, cols]`` (x = p0 + p1 + p2 to the last fp32 ulp)."""
    global launch_count
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    rows, cols = x.shape
    out = torch.empty(3, rows, cols, dtype=torch.bfloat16, device=x.device)
    _check(load().b200_split_f32(out.data_ptr(), x.data_ptr(), rows, cols, x.stride(0), _stream()), "b200_split_f32")
    launch_count += 1
    return out


def gemm_f32(A3: torch.Tensor, a_major: int, B3: torch.Tensor, b_major: int, out: torch.Tensor, M: int, N: int, K: int,
             epilogue: int = EPI_STORE, aux: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None,
             bias: Optional[torch.Tensor] = None, alpha: float = 1.0, max_ctas: int = 0) -> torch.Tensor:
    """fp32 ``out[M,N] = epilogue(A . B)`` from the split planes of A and B (``split3``): six bf16 tensor-core products
    accumulated in fp32 (csrc/gemm_f32.cu).  Epilogues STORE, RELU, GELU, RELU_BWD, GELU_BWD, ADD, SGD; all tensors fp32."""
    global launch_count
    assert A3.dtype == torch.bfloat16 and B3.dtype == torch.bfloat16 and A3.dim() == 3 and B3.dim() == 3 and out.dtype == torch.float32
    rc = load().b200_gemm_f32(A3.data_ptr(), A3.shape[2], a_major, B3.data_ptr(), B3.shape[2], b_major, out.data_ptr(), out.stride(0),
                              M, N, K, epilogue, _ptr(aux), aux.stride(0) if aux is not None else 0, _ptr(out2),
                              out2.stride(0) if out2 is not None else 0, _ptr(bias), float(alpha), int(max_ctas), _stream())
    _check(rc, "b200_gemm_f32")
    launch_count += 1
    return out


# --------------------------------------------------------------------------------------------- elementwise
def sgd_(param: torch.Tensor, grad: torch.Tensor, lr: float) -> None:
    global launch_count
    assert param.dtype == torch.bfloat16 and param.is_contiguous() and grad.is_contiguous()
    _check(load().b200_sgd_bf16(param.data_ptr(), grad.data_ptr(), param.numel(), float(lr),
                                int(grad.dtype == torch.float32), _stream()), "b200_sgd_bf16")
    launch_count += 1


def adam_(param: torch.Tensor, grad: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, beta1: float,
          beta2: float, eps: float, step: int) -> None:
    global launch_count
    assert param.dtype == torch.bfloat16 and m.dtype == torch.float32 and v.dtype == torch.float32
    _check(load().b200_adam_bf16(param.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), param.numel(),
                                 float(lr), float(beta1), float(beta2), float(eps), int(step),
                                 int(grad.dtype == torch.float32), _stream()), "b200_adam_bf16")
    launch_count += 1


def sgd_f32_(param: torch.Tensor, grad: torch.Tensor, lr: float) -> None:
    global launch_count
    assert param.dtype == torch.float32 and grad.dtype == torch.float32 and param.is_contiguous() and grad.is_contiguous()
    _check(load().b200_sgd_f32(param.data_ptr(), grad.data_ptr(), param.numel(), float(lr), _stream()), "b200_sgd_f32")
    launch_count += 1


def adam_f32_(param: torch.Tensor, grad: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, beta1: float, beta2: float,
              eps: float, step: int) -> None:
    global launch_count
    assert param.dtype == torch.float32 and grad.dtype == torch.float32 and m.dtype == torch.float32 and v.dtype == torch.float32
    _check(load().b200_adam_f32(param.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), param.numel(), float(lr),
                                float(beta1), float(beta2), float(eps), int(step), _stream()), "b200_adam_f32")
    launch_count += 1


_colsum_scratch: dict = {}


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """``out[n] (+)= sum_t x[t, n]`` (fp32) for a bf16 ``[T, N]`` matrix - the bias gradient."""
    global launch_count
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    T, N = x.shape
    lib = load()
    need = int(lib.b200_colsum_scratch_bytes(T, N))
    key = x.device.index or 0
    sc = _colsum_scratch.get(key)
    if sc is None or sc.numel() < need:
        sc = _colsum_scratch[key] = torch.empty(max(need, 1 << 20), dtype=torch.uint8, device=x.device)
    if out is None:
        out = torch.empty(N, dtype=torch.float32, device=x.device)
        accumulate = False
    _check(lib.b200_colsum_bf16(out.data_ptr(), x.data_ptr(), x.stride(0), T, N, int(accumulate), sc.data_ptr(), sc.numel(), _stream()),
           "b200_colsum_bf16")
    launch_count += 2
    return out


def bias_rows_(out: torch.Tensor, bias: Optional[torch.Tensor], residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out[t, :] = bias (+ residual[t, :])`` (bf16 out / residual, fp32 bias)."""
    global launch_count
    T, N = out.shape
    _check(load().b200_bias_rows_bf16(out.data_ptr(), out.stride(0), _ptr(residual), residual.stride(0) if residual is not None else 0,
                                      _ptr(bias), T, N, _stream()), "b200_bias_rows_bf16")
    launch_count += 1
    return out


def transpose(x: torch.Tensor) -> torch.Tensor:
    """Contiguous ``x.T`` of a bf16 matrix on the native tiled-transpose kernel."""
    global launch_count
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    R, Cc = x.shape
    assert R % 2 == 0 and x.stride(0) % 2 == 0, "transpose: even row count and pitch (4-byte paired accesses)"
    out = torch.empty(Cc, R, dtype=torch.bfloat16, device=x.device)
    _check(load().b200_transpose_bf16(out.data_ptr(), out.stride(0), x.data_ptr(), x.stride(0), R, Cc, _stream()), "b200_transpose_bf16")
    launch_count += 1
    return out


def relu(x: torch.Tensor) -> torch.Tensor:
    global launch_count
    out = torch.empty_like(x)
    _check(load().b200_relu_bf16(out.data_ptr(), x.data_ptr(), x.numel(), _stream()), "b200_relu_bf16")
    launch_count += 1
    return out


def relu_bwd_(dy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    global launch_count
    _check(load().b200_relu_bwd_bf16(dy.data_ptr(), x.data_ptr(), x.numel(), _stream()), "b200_relu_bwd_bf16")
    launch_count += 1
    return dy


def add(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    global launch_count
    out = torch.empty_like(a) if out is None else out
    _check(load().b200_add_bf16(out.data_ptr(), a.data_ptr(), b.data_ptr(), a.numel(), _stream()), "b200_add_bf16")
    launch_count += 1
    return out


def randn_(out: torch.Tensor, seed: int, scale: float = 1.0) -> torch.Tensor:
    global launch_count
    assert out.dtype == torch.bfloat16 and out.is_contiguous()
    _check(load().b200_randn_bf16(out.data_ptr(), out.numel(), int(seed), float(scale), _stream()), "b200_randn_bf16")
    launch_count += 1
    return out


def cast_to_bf16(x: torch.Tensor) -> torch.Tensor:
    global launch_count
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    _check(load().b200_cast_f32_to_bf16(out.data_ptr(), x.data_ptr(), x.numel(), _stream()), "b200_cast_f32_to_bf16")
    launch_count += 1
    return out


def cast_to_f32(x: torch.Tensor) -> torch.Tensor:
    global launch_count
    out = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    _check(load().b200_cast_bf16_to_f32(out.data_ptr(), x.data_ptr(), x.numel(), _stream()), "b200_cast_bf16_to_f32")
    launch_count += 1
    return out


_sumsq_scratch: dict[int, torch.Tensor] = {}


def sumsq(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Deterministic sum of squares (fp32 scalar tensor on the device)."""
    global launch_count
    lib = load()
    dev = x.device.index or 0
    sc = _sumsq_scratch.get(dev)
    if sc is None:
        sc = torch.zeros(int(lib.b200_sumsq_scratch_bytes()), dtype=torch.uint8, device=x.device)
        _sumsq_scratch[dev] = sc
    out = torch.empty(1, dtype=torch.float32, device=x.device) if out is None else out
    _check(lib.b200_sumsq_bf16(out.data_ptr(), x.data_ptr(), x.numel(), sc.data_ptr(), sc.numel(), _stream()),
           "b200_sumsq_bf16")
    launch_count += 1
    return out


def sumsq_static(x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """Sum of squares into the caller's ``out`` (1 fp32 element) without allocating anything: no scratch tensor, no
    output tensor, no library-side allocation (``b200_sumsq_bf16_static``).  Calls on one device must not overlap."""
    global launch_count
    _check(load().b200_sumsq_bf16_static(out.data_ptr(), x.data_ptr(), x.numel(), _stream()), "b200_sumsq_bf16_static")
    launch_count += 1
    return out


def set_update_rounding(seed: int) -> None:
    """0 = round optimizer updates to nearest; else stochastic rounding with this seed (see include/b200ffn.h)."""
    load().b200_set_update_rounding(int(seed) & 0xFFFFFFFF)


def fill_zero_(buf: torch.Tensor) -> None:
    global launch_count
    _check(load().b200_fill_zero(buf.data_ptr(), buf.numel() * buf.element_size(), _stream()), "b200_fill_zero")
    launch_count += 1
