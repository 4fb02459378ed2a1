"""Public training API: one ``Engine`` per rank.

    eng = Engine(cfg, method="ddp")            # under torchrun, or world 1
    for metric in eng.train(host_batches):     # pinned host -> device staging, step, metric read-back
        ...

``Engine.step(x, dloss_dx)`` runs one optimizer step on device tensors (what ``bench.py`` device-times);
``Engine.train`` is the end-to-end path a user calls: every step copies its inputs from pinned host memory on a
side stream (overlapped with the previous step's compute), runs the step, reduces the block output to a scalar
metric on the device and reads it back (one step late, so the read never stalls the GPU).
"""
from __future__ import annotations

from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

import torch

from .config import INIT_SCALE, TrainConfig
from .models.ffn import Params
from .ops import get_ops
from .parallel.api import build_trainer
from .parallel.comm import LocalComm, make_comm
from .parallel.launcher import init_from_env
from .utils.data import PinnedPrefetcher, ShardedPrefetcher


def init_params_on_device(cfg: TrainConfig, device: torch.device, seed: int = 1234,
                          scale: float = INIT_SCALE) -> List[Params]:
    """Random-init weights of the configured architecture, generated on the device (same seed => the same
    replica on every rank).  Layout/scale as ``train_ffns.py:35-39``."""
    dtype = cfg.torch_dtype
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = []
    for _ in range(cfg.layers):
        w1 = (scale * torch.randn((cfg.ffn_dim, cfg.model_size), generator=g, device=device)).to(dtype)
        w2 = (scale * torch.randn((cfg.model_size, cfg.ffn_dim), generator=g, device=device)).to(dtype)
        lp = [w1, w2]
        if cfg.bias:
            lp += [torch.zeros(cfg.ffn_dim, dtype=torch.float32, device=device), torch.zeros(cfg.model_size, dtype=torch.float32, device=device)]
        out.append(lp)
    return out


class Engine:
    def __init__(self, cfg: TrainConfig, method: str = "ddp", layers_params: Optional[Sequence[Params]] = None,
                 rank: Optional[int] = None, world: Optional[int] = None, device: Optional[torch.device] = None,
                 seed: int = 1234):
        if rank is None or world is None or device is None:
            rank, world, device = init_from_env(cfg.device, cfg.backend)
        self.cfg, self.method, self.rank, self.world, self.device = cfg, method, rank, world, device
        self.ops = get_ops(cfg.impl, device)
        self.comm = make_comm() if world > 1 else LocalComm()
        if layers_params is None:
            layers_params = init_params_on_device(cfg, device, seed)
        self.trainer = build_trainer(method, self.ops, self.comm, cfg, layers_params, device)
        del layers_params
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._graphs: dict = {}

    # ---- device-side step ---------------------------------------------------------------------------------
    def step(self, x: torch.Tensor, dloss_dx: torch.Tensor) -> torch.Tensor:
        if self.cfg.cuda_graph and x.is_cuda:
            g = self._graphs.get((tuple(x.shape), x.dtype))
            if g is None:                      # first step of this shape: warm up, capture (weights are put back)
                from .utils.graph import GraphedStep
                g = self._graphs[(tuple(x.shape), x.dtype)] = GraphedStep(self.trainer, x, dloss_dx)
            return g.step(x, dloss_dx)
        return self.trainer.step(x, dloss_dx)

    @property
    def tokens_per_step_global(self) -> int:
        """DDP/FSDP consume ``world`` batches per optimizer step; 1 GPU and MP one (``train_ffns.py:182,:324``)."""
        return self.cfg.tokens * (self.world if self.method in ("ddp", "fsdp") else 1)

    # ---- end-to-end loop ----------------------------------------------------------------------------------
    def train(self, host_batches: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> Iterator[float]:
        """Yield one scalar metric (sum of squares of the stack output) per step, read back from the device."""
        if self.device.type != "cuda":
            for x, dy in host_batches:
                y = self.step(x.to(self.cfg.torch_dtype), dy.to(self.cfg.torch_dtype))
                yield float((y.float() ** 2).sum())
            return
        rows = None
        if self.method == "mp" and self.world > 1 and self.cfg.mp_mode in ("sp", "spf"):
            # sequence-parallel MP reads only its own token shard of the batch: stage 1/world of the bytes
            ts = self.cfg.tokens // self.world
            rows = slice(self.rank * ts, (self.rank + 1) * ts)
        sw = getattr(self.trainer, "sw", None)
        if self.method == "mp" and self.world > 1 and rows is None and sw is not None and self.cfg.mp_input == "sharded" \
                and self.cfg.tokens % self.world == 0:
            # all-reduce form: every rank needs the whole batch - stage 1/world of it over PCIe, gather the rest over NVLink
            pf = ShardedPrefetcher(host_batches, self.device, sw, self.rank, self.world)
        else:
            pf = PinnedPrefetcher(host_batches, self.device, rows=rows)
        pending = None
        host = torch.empty(2, dtype=torch.float32).pin_memory()
        slot = 0
        for x, dy in pf:
            self.h2d_bytes += pf.last_bytes
            y = self.step(x, dy)
            m = self.ops.sumsq(y)
            host[slot:slot + 1].copy_(m, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            self.d2h_bytes += 4
            if pending is not None:
                pev, pslot = pending
                pev.synchronize()
                yield float(host[pslot])
            pending = (ev, slot)
            slot ^= 1
        if pending is not None:
            pending[0].synchronize()
            yield float(host[pending[1]])

    def full_params(self) -> List[Params]:
        return self.trainer.full_params()

    def close(self) -> None:
        """Release the trainer and, with it, the symmetric memory of the fused paths (collective under > 1 rank), so that
        another ``Engine`` can be built in the same process (bench.py measures DDP, FSDP and MP back to back)."""
        import gc
        self.trainer = None
        self.comm = None
        self._graphs = {}
        gc.collect()
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        if self.device.type == "cuda":
            from .parallel.symm import release_world
            release_world()                    # no-op when no symmetric world was created
        if self.device.type == "cuda":
            torch.cuda.empty_cache()
