"""Op backends: ``get_ops("native")`` -> sm_100a kernels, ``get_ops("torch")`` -> PyTorch (CPU/gloo + oracle)."""
from __future__ import annotations

import torch

from .torch_ops import TorchOps


def get_ops(impl: str = "auto", device: torch.device | str | None = None, max_ctas: int = 0):
    """``auto`` picks the native kernels on CUDA and PyTorch on CPU.  Asking for ``native`` on a box without
    CUDA, or with a library that fails to load, raises - there is no silent fallback."""
    dev = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
    if impl == "auto":
        impl = "native" if dev.type == "cuda" else "torch"
    if impl == "torch":
        return TorchOps()
    if impl == "native":
        if dev.type != "cuda":
            raise RuntimeError("impl=native needs a CUDA device (sm_100a)")
        from .cuda_ops import CudaOps
        return CudaOps(max_ctas=max_ctas)
    raise ValueError(f"unknown impl {impl!r}")
