"""Strategy drivers with the reference's uniform signature

    fn(layers_params, seeds, batch_size, model_size) -> layers_params

(``train_1gpu`` ``train_ffns.py:101``, ``train_ddp`` ``:174``, ``train_fsdp`` ``:262``, ``train_tp`` ``:315``;
``batch_size`` is the token count per step, ``:379``).  Each driver spawns one process per rank, lets every rank
build ITS OWN replica/shard from the shared CPU parameters (no parent-side CUDA contexts), runs the
per-rank trainer and writes the result back into CPU shared-memory tensors owned by the parent.
"""
from __future__ import annotations

import copy
from typing import List, Optional, Sequence

import torch

from ..config import TrainConfig
from ..models.ffn import Params
from ..ops import get_ops
from ..utils.data import device_mock_data, mock_data, split_seeds
from ..utils.profiler import enable_ranges, profile_if
from ..utils.timing import DeviceTimer, max_over_ranks, tokens_per_second
from .base import fsdp_shard, place, tp_shard, tp_shard_dim
from .comm import LocalComm, make_comm
from .ddp import DDPTrainer
from .fsdp import FSDPTrainer
from .launcher import launch, pick_backend, pick_device, visible_world
from .mp import MPTrainer
from .single import SingleTrainer

METHODS = ("single", "ddp", "fsdp", "mp")


def build_trainer(method: str, ops, comm, cfg: TrainConfig, layers_params: Sequence[Params], device: torch.device):
    """Place/shard the full CPU parameters for this rank and return its trainer."""
    dtype = cfg.torch_dtype
    r, n = comm.rank, comm.world
    if cfg.comm == "native" and n > 1:
        # unfused trainers (so Adam / bias / residual / SP all work) on the framework's own peer-memory collectives
        if device.type != "cuda" or dtype != torch.bfloat16:
            raise RuntimeError("comm=native needs CUDA and bf16 parameters (the native collectives sum bf16 / fp32)")
        from .symm import SymmComm, get_symm_world
        full_bytes = 2 * max(cfg.model_size * cfg.ffn_dim, cfg.tokens * max(cfg.model_size, cfg.ffn_dim // n))
        comm = SymmComm(get_symm_world(comm, device), staging_bytes=max(64 << 20, full_bytes + 4096), group=getattr(comm, "group", None))
    if method == "single" or n == 1:        # a world of one: every strategy degenerates to the single-device loop
        return SingleTrainer(ops, cfg, place(layers_params, device, dtype))
    if method == "ddp":
        if cfg.comm in ("auto", "fused") and _fused_available(ops, comm, cfg, "ddp"):
            from .fused import FusedDDPTrainer
            return FusedDDPTrainer(ops, comm, cfg, layers_params, device)
        return DDPTrainer(ops, comm, cfg, place(layers_params, device, dtype))
    if method == "fsdp":
        shards = [[fsdp_shard(p, r, n) for p in lp] for lp in layers_params]
        if cfg.comm in ("auto", "fused") and _fused_available(ops, comm, cfg, "fsdp"):
            if cfg.fsdp_gather == "fused" and cfg.tokens % 8 == 0 and not cfg.bias:   # (transposed activations: a per-row bias has no epilogue yet)
                # parameter all-gather inside the consuming GEMMs, forward and backward
                from .fused_fsdp_t import FusedFSDPGatherTrainer
                return FusedFSDPGatherTrainer(ops, comm, cfg, place(shards, device, dtype), device)
            from .fused import FusedFSDPTrainer
            return FusedFSDPTrainer(ops, comm, cfg, place(shards, device, dtype), device)
        return FSDPTrainer(ops, comm, cfg, place(shards, device, dtype))
    if method == "mp":
        shards = [[tp_shard(p, j, r, n) for j, p in enumerate(lp)] for lp in layers_params]
        if cfg.comm in ("auto", "fused") and _fused_available(ops, comm, cfg, "mp"):
            from .fused import FusedMPTrainer
            return FusedMPTrainer(ops, comm, cfg, place(shards, device, dtype), device)
        return MPTrainer(ops, comm, cfg, place(shards, device, dtype))
    raise ValueError(f"unknown method {method!r}")


def _fused_available(ops, comm, cfg: TrainConfig, method: str = "ddp") -> bool:
    """The peer-memory fused kernels need the native ops, bf16, >1 rank and TMA-legal shapes (FSDP additionally needs
    every row shard to be a whole number of 128-row tiles).  SGD or Adam, ReLU or GELU, with or without bias / residual."""
    # Adam: inside the kernel for DDP / FSDP (the tile owner applies it and owns the moments); MP: fused collectives + the
    # native Adam kernel on the shard-local dW.
    ok = (getattr(ops, "name", "") == "native" and comm.world > 1 and cfg.dtype in ("bf16", "bfloat16")
          and cfg.model_size % 8 == 0 and cfg.ffn_dim % 8 == 0)
    if method == "fsdp":
        ok = ok and cfg.model_size % (128 * comm.world) == 0 and cfg.ffn_dim % (128 * comm.world) == 0
    if method == "mp":
        # (a residual rides on the kernels' in-place update mode: output pre-filled with the residual, alpha = -1)
        ok = ok and (cfg.ffn_dim // comm.world) % 8 == 0
        if cfg.mp_mode in ("sp", "spf"):
            ok = ok and cfg.tokens % (128 * comm.world) == 0
        if cfg.mp_mode == "spf":
            ok = ok and cfg.model_size % 64 == 0
    if not ok:
        if cfg.comm == "fused":
            raise RuntimeError("comm=fused needs impl=native, bf16, world>1, dims % 8 == 0 "
                               "(FSDP: dims % (128*world) == 0; MP sp forms: tokens % (128*world) == 0)")
        return False
    try:
        from . import fused  # noqa: F401
        return fused.available()
    except ImportError:
        if cfg.comm == "fused":
            raise
        return False


def _write_back(method: str, trainer, comm, out_params: Sequence[Params]) -> None:
    """Copy this rank's result into the parent's shared CPU tensors (full layout, any strategy)."""
    r, n = comm.rank, comm.world
    for i, lp in enumerate(trainer.params if method in ("single", "ddp") else trainer.shards):
        for j, p in enumerate(lp):
            dst = out_params[i][j]
            src = p.detach().to("cpu", dtype=dst.dtype)
            if method in ("single", "ddp"):
                if r == 0:
                    dst.copy_(src)
            elif method == "fsdp":
                dst.chunk(n, dim=0)[r].copy_(src)
            else:
                dim = tp_shard_dim(j)
                if dim is None:
                    if r == 0:
                        dst.copy_(src)
                else:
                    dst.chunk(n, dim=dim)[r].copy_(src)


def _rank_main(rank: int, world: int, device: torch.device, method: str, cfg_dict: dict,
               layers_params: Sequence[Params], seeds: torch.Tensor, tokens: int, model_size: int,
               out_params: Sequence[Params], stats: torch.Tensor, opt_io: Optional[dict] = None) -> None:
    cfg = TrainConfig(**cfg_dict)
    ops = get_ops(cfg.impl, device)
    comm = make_comm() if world > 1 else LocalComm()
    trainer = build_trainer(method, ops, comm, cfg, layers_params, device)
    if opt_io is not None and opt_io.get("t_in", 0) > 0:             # resume: Adam moments from the checkpoint's full layout
        from .opt_state import import_state
        import_state(method if world > 1 else "single", trainer, comm, opt_io["m_in"], opt_io["v_in"], opt_io["t_in"])
    my_seeds = split_seeds(seeds, world)[rank] if method in ("ddp", "fsdp") else seeds
    if cfg.data == "device" and device.type == "cuda":
        batches = device_mock_data(my_seeds, tokens, model_size, cfg.torch_dtype, device)
    else:
        batches = mock_data(my_seeds, tokens, model_size, cfg.torch_dtype, device)
    seeds_per_step = world if method in ("ddp", "fsdp") else 1
    step = trainer.step
    if cfg.cuda_graph and world == 1 and device.type == "cuda":
        from ..utils.graph import GraphedStep
        graphs: dict = {}

        def step(x, dy):                                             # noqa: F811 - graph replay of the same step
            g = graphs.get(tuple(x.shape))
            if g is None:
                g = graphs[tuple(x.shape)] = GraphedStep(trainer, x, dy)
            return g.step(x, dy)
    if cfg.profile:
        enable_ranges(True)                                          # per-block NVTX ranges in the model loops
    with profile_if(cfg.profile, rank):
        with DeviceTimer(device) as t:
            if cfg.save_every and cfg.save_path:
                for k, (x, dy) in enumerate(batches):
                    step(x, dy)
                    if (k + 1) % cfg.save_every == 0:                # periodic checkpoint: full, unsharded layout
                        full = trainer.full_params()                 # collective on every rank
                        opt_sd = {"kind": cfg.optimizer, "t": k + 1}
                        if opt_io is not None:                       # Adam: every rank exports its pieces, then rank 0 saves
                            from .opt_state import export_state
                            opt_sd["t"] = export_state(method if world > 1 else "single", trainer, comm, opt_io["m_out"], opt_io["v_out"])
                            comm.barrier()                           # all pieces are in the shared tensors
                            opt_sd.update(m=[[t_.clone() for t_ in l] for l in opt_io["m_out"]],
                                          v=[[t_.clone() for t_ in l] for l in opt_io["v_out"]])
                        if rank == 0:
                            from ..utils.checkpoint import save_checkpoint
                            save_checkpoint(f"{cfg.save_path}.step{(k + 1) * seeds_per_step}", full, step=(k + 1) * seeds_per_step,
                                            args=cfg.to_dict(), seeds=seeds, optimizer=opt_sd)
                        if opt_io is not None:
                            comm.barrier()                           # nobody overwrites the shared tensors while rank 0 reads them
            elif step is not trainer.step:
                for x, dy in batches:
                    step(x, dy)
            else:
                trainer.run(batches)
    ms = max_over_ranks(t.ms, comm, device)
    steps = len(my_seeds)
    global_tokens = tokens * seeds_per_step
    if rank == 0:
        stats[0] = ms
        stats[1] = tokens_per_second(global_tokens, steps, ms)
        # model FLOPs: 6 GEMMs of 2*T*D*F per block-step (the reference executes a 7th, recomputing h, train_ffns.py:63)
        stats[2] = 12.0 * global_tokens * cfg.model_size * cfg.ffn_dim * cfg.layers * steps / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
        stats[3] = torch.cuda.max_memory_allocated(device) / 2 ** 30 if device.type == "cuda" else 0.0
    if opt_io is not None:
        from .opt_state import export_state
        t_done = export_state(method if world > 1 else "single", trainer, comm, opt_io["m_out"], opt_io["v_out"])
        if rank == 0:
            stats[4] = t_done
    _write_back(method if world > 1 else "single", trainer, comm, out_params)


def _run(method: str, layers_params: Sequence[Params], seeds: torch.Tensor, tokens: int, model_size: int,
         cfg: Optional[TrainConfig], opt_state: Optional[dict] = None) -> List[Params]:
    cfg = copy.deepcopy(cfg) if cfg is not None else TrainConfig(model_size=model_size)
    cfg.model_size = model_size
    cfg.batch_size, cfg.seq_len = tokens, 1          # the drivers take the token count directly (train_ffns.py:379)
    cfg.ffn_dim = layers_params[0][0].shape[0]
    cfg.bias = len(layers_params[0]) == 4
    device = pick_device(cfg.device)
    world = 1 if method == "single" else visible_world(cfg.world_size, device)
    backend = pick_backend(cfg.backend, device)
    if method in ("ddp", "fsdp") and len(seeds) % world:
        raise AssertionError(f"len(seeds)={len(seeds)} must be divisible by the world size {world}")   # :175, :263
    shared_in = [[p.detach().clone().share_memory_() for p in lp] for lp in layers_params]
    out = [[torch.empty_like(p).share_memory_() for p in lp] for lp in layers_params]
    stats = torch.zeros(5, dtype=torch.float64).share_memory_()
    opt_io = None
    if cfg.optimizer == "adam":          # optimizer state travels in the parameters' full layout (parallel/opt_state.py)
        from .opt_state import empty_like_params
        opt_io = {"m_out": empty_like_params(layers_params), "v_out": empty_like_params(layers_params), "t_in": 0}
        if opt_state and opt_state.get("kind") == "adam" and opt_state.get("t", 0) > 0:
            opt_io["m_in"] = [[t.detach().float().clone().share_memory_() for t in l] for l in opt_state["m"]]
            opt_io["v_in"] = [[t.detach().float().clone().share_memory_() for t in l] for l in opt_state["v"]]
            opt_io["t_in"] = int(opt_state["t"])
    args = (method, cfg.to_dict(), shared_in, seeds, tokens, model_size, out, stats, opt_io)
    if world == 1:
        dev = torch.device("cuda", 0) if device == "cuda" else torch.device("cpu")
        _rank_main(0, 1, dev, *args)
    else:
        launch(world, _rank_main, args, backend=backend, device=device, port=cfg.master_port if cfg.master_port != 29500 else 0)
    _run.last_stats = {"ms": float(stats[0]), "tokens_per_s": float(stats[1]), "model_tflops": float(stats[2]),
                       "max_mem_gib": float(stats[3]), "world": world}
    _run.last_opt = {"kind": cfg.optimizer, "t": int(stats[4])}
    if opt_io is not None:
        _run.last_opt.update(m=[list(l) for l in opt_io["m_out"]], v=[list(l) for l in opt_io["v_out"]])
    return [list(lp) for lp in out]


_run.last_stats = {}
_run.last_opt = {}


def last_optimizer_state() -> dict:
    """Optimizer state of the most recent driver call in the checkpoint's full layout: ``{"kind", "t"}`` plus, for Adam,
    ``m`` / ``v`` shaped like the parameters.  Pass it back as ``opt_state=`` (any strategy) to resume."""
    return dict(_run.last_opt)


def last_stats() -> dict:
    """Device-timed milliseconds (max over ranks) and tokens/s of the most recent driver call."""
    return dict(_run.last_stats)


def train_1gpu(layers_params, seeds, batch_size, model_size, cfg: Optional[TrainConfig] = None, opt_state: Optional[dict] = None):
    return _run("single", layers_params, seeds, batch_size, model_size, cfg, opt_state)


def train_ddp(layers_params, seeds, batch_size, model_size, cfg: Optional[TrainConfig] = None, opt_state: Optional[dict] = None):
    return _run("ddp", layers_params, seeds, batch_size, model_size, cfg, opt_state)


def train_fsdp(layers_params, seeds, batch_size, model_size, cfg: Optional[TrainConfig] = None, opt_state: Optional[dict] = None):
    return _run("fsdp", layers_params, seeds, batch_size, model_size, cfg, opt_state)


def train_tp(layers_params, seeds, batch_size, model_size, cfg: Optional[TrainConfig] = None, opt_state: Optional[dict] = None):
    return _run("mp", layers_params, seeds, batch_size, model_size, cfg, opt_state)
