"""CUDA-graph replay of a whole optimizer step (launch-bound configurations).

A step of a small stack is a few dozen kernels of a few microseconds each: the host's launch rate, not the GPU, sets
the step time.  ``GraphedStep`` captures ``trainer.step`` once - native tcgen05 GEMMs, elementwise kernels and the
torch glue between them all land on the capture stream - and replays it with one ``cudaGraphLaunch`` per step.

Scope: one GPU, SGD.  Not capturable: Adam (its step count is a kernel ARGUMENT, frozen at capture), and the
peer-memory fused kernels (their per-call sequence numbers are kernel arguments too, and they wait on other ranks).

Validated on B200 in round 2 (``tests/test_train_gpu.py::test_cuda_graph_replay_matches_eager_steps``).

The reference has no counterpart (eager PyTorch, ``train_ffns.py:101-116``).
"""
from __future__ import annotations

from typing import List

import torch


class GraphedStep:
    def __init__(self, trainer, x_like: torch.Tensor, dy_like: torch.Tensor, warmup: int = 2):
        if not x_like.is_cuda:
            raise ValueError("CUDA graphs need a CUDA device")
        opt = getattr(trainer, "opt", None)
        if opt is not None and opt.kind != "sgd":
            raise ValueError("GraphedStep supports SGD only (Adam's step count is a kernel argument)")
        if getattr(getattr(trainer, "comm", None), "world", 1) > 1:
            raise ValueError("GraphedStep is single-GPU: the fused multi-GPU kernels carry per-call sequence numbers")
        self.trainer = trainer
        self.x = torch.empty_like(x_like)
        self.dy = torch.empty_like(dy_like)
        self.x.copy_(x_like)
        self.dy.copy_(dy_like)
        params: List[torch.Tensor] = [p for lp in trainer.full_params() for p in lp]
        keep = [p.clone() for p in params]          # warm-up steps really train: put the weights back afterwards
        side = torch.cuda.Stream(device=x_like.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            # first calls set function attributes / allocate the error slot: none of that may happen during capture
            for _ in range(max(1, warmup)):
                trainer.step(self.x, self.dy)
        torch.cuda.current_stream().wait_stream(side)
        for p, k in zip(params, keep):
            p.copy_(k)
        del keep
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.y = trainer.step(self.x, self.dy)
        self.replays = 0

    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        """Same contract as ``trainer.step``; the returned tensor is the graph's static output buffer (overwritten by
        the next replay)."""
        self.x.copy_(x, non_blocking=True)
        self.dy.copy_(dloss_dx, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        return self.y
