"""Device-side timing: CUDA events on the launching stream, max over ranks (never host wall-clock for a metric;
the reference times with ``time.time()`` around spawn + NCCL init + CPU RNG + H2D, ``train_ffns.py:378-380``)."""
from __future__ import annotations

import time

import torch


class DeviceTimer:
    """``with DeviceTimer(device) as t: ...`` then ``t.ms`` (CUDA events on GPU, perf_counter on CPU)."""

    def __init__(self, device: torch.device):
        self.device = device
        self.ms = 0.0

    def __enter__(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
            self._s = torch.cuda.Event(enable_timing=True)
            self._e = torch.cuda.Event(enable_timing=True)
            self._s.record()
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.device.type == "cuda":
            self._e.record()
            torch.cuda.synchronize(self.device)
            self.ms = self._s.elapsed_time(self._e)
        else:
            self.ms = (time.perf_counter() - self._t0) * 1e3
        return False


def max_over_ranks(ms: float, comm, device: torch.device) -> float:
    t = torch.tensor([ms], dtype=torch.float64, device=device if device.type == "cuda" else "cpu")
    if t.is_cuda:
        t = t.float()
    return float(comm.all_reduce_max(t).item())


def tokens_per_second(tokens_per_step_global: int, steps: int, ms: float) -> float:
    return tokens_per_step_global * steps / (ms * 1e-3) if ms > 0 else float("nan")
