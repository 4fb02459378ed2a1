"""Profiling hooks.  ``torch_profile_rank_0`` mirrors the reference decorator (``train_ffns.py:129-141``):
wraps a per-rank function in ``torch.profiler`` and has rank 0 export ``trace_profiler_trace.json``.  NVTX ranges
mark phases for ncu/nsys captures."""
from __future__ import annotations

import contextlib
import functools
import os

import torch

TRACE_FILE = "trace_profiler_trace.json"


def torch_profile_rank_0(func):
    @functools.wraps(func)          # keeps the wrapped function picklable by qualified name under spawn
    def wrapper(*args, **kwargs):
        from torch.profiler import ProfilerActivity, profile
        rank = args[0]
        acts = [ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if torch.cuda.is_available() else [])
        with profile(activities=acts, record_shapes=True, with_stack=True) as prof:
            out = func(*args, **kwargs)
        if rank == 0:
            prof.export_chrome_trace(TRACE_FILE)
            print("Profiler exported")
        return out
    return wrapper


@contextlib.contextmanager
def profile_if(enabled: bool, rank: int, path: str = TRACE_FILE):
    if not enabled:
        yield None
        return
    from torch.profiler import ProfilerActivity, profile
    acts = [ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if torch.cuda.is_available() else [])
    with profile(activities=acts, record_shapes=True, with_stack=True) as prof:
        yield prof
    if rank == 0:
        prof.export_chrome_trace(path)
        print("Profiler exported")


@contextlib.contextmanager
def nvtx(name: str):
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


# Per-block ranges inside the model loops (``ffn_fwd`` / ``ffn_bwd``): off by default - one shared null context, no
# per-call allocation - and switched on by ``--profile`` or ``B200_NVTX=1`` so timeline captures show the block
# structure above the kernel names.
_NULL = contextlib.nullcontext()
_ranges_on = os.environ.get("B200_NVTX") == "1"


def enable_ranges(on: bool = True) -> None:
    global _ranges_on
    _ranges_on = bool(on)


def block_range(name: str):
    return nvtx(name) if _ranges_on else _NULL
