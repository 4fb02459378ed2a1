"""Run configuration and the reference's module constants.

Reference: ``train_ffns.py:29-30`` (``LR``, ``DLOSS_DX_COEF``), ``:35`` (init scale), ``:344-350`` (CLI flags).
"""
from __future__ import annotations

from dataclasses import asdict, dataclass, field
from typing import Optional

import torch

LR = 0.00001            # train_ffns.py:29
DLOSS_DX_COEF = 0.1     # train_ffns.py:30
INIT_SCALE = 2e-2       # train_ffns.py:35
SEED_RANGE = 100_000    # train_ffns.py:360

METHOD_NAMES = {1: "train_1gpu", 2: "train_ddp", 3: "train_fsdp", 4: "train_tp"}   # train_ffns.py:373

_DTYPES = {"fp32": torch.float32, "float32": torch.float32, "bf16": torch.bfloat16, "bfloat16": torch.bfloat16}


def parse_dtype(name: str) -> torch.dtype:
    try:
        return _DTYPES[name.lower()]
    except KeyError as e:
        raise ValueError(f"unknown dtype {name!r}; expected one of {sorted(_DTYPES)}") from e


@dataclass
class TrainConfig:
    """Everything a strategy needs.  The first seven fields are the reference's flags with its defaults."""
    num_steps: int = 1
    batch_size: int = 8
    seq_len: int = 1024
    layers: int = 1
    model_size: int = 4
    method: int = 0
    random_seed: int = 0
    # ---- extensions ----
    ffn_dim: Optional[int] = None          # default 4 * model_size (train_ffns.py:361)
    dtype: str = "fp32"                    # reference is fp32; bf16 is the performance target
    activation: str = "relu"               # reference: ReLU (:47); north-star: GELU
    residual: bool = False                 # reference: none (:54-58)
    bias: bool = False                     # reference: none (:35)
    optimizer: str = "sgd"                 # reference: SGD (:114); extension: adam
    lr: float = LR
    grad_scale: float = 1.0                # reference sums gradients across ranks, no 1/world (:165)
    world_size: int = 0                    # 0 = all visible GPUs (reference behaviour, :25)
    backend: str = "auto"                  # auto | nccl | gloo
    device: str = "auto"                   # auto | cuda | cpu
    impl: str = "auto"                     # auto | native (sm_100a kernels) | torch (cuBLAS/ATen fallback + oracle)
    comm: str = "auto"                     # auto | fused (collective inside the GEMM) | native (peer-memory collectives) | nccl
    recompute: bool = False                # reference recomputes h in backward (:63); we keep it by default
    mp_mode: str = "ar"                    # ar = reference all-reduce form (:303,:309); sp = AG/RS sequence-parallel; spf = sp with the AG inside the GEMM
    fsdp_gather: str = "fused"               # fused FSDP parameter all-gather: fused = inside the consuming GEMMs, forward and backward
                                           # (parallel/fused_fsdp_t.py; default: ties the prefetch at 8 GPUs, 325.4 vs 325.0 ms/step);
                                           # dma = copy-engine prefetch one layer ahead (:236-249); biases need dma (chosen automatically)
    mp_input: str = "sharded"                 # Engine.train under MP all-reduce form: sharded = 1/world over PCIe + NVLink all-gather
                                           # (ShardedPrefetcher; default: 87.4 vs 96.6 ms/step end to end at 8 GPUs); full = every
                                           # rank stages the whole batch like the reference (:294, :324)
    stochastic_rounding: bool = False      # bf16 weights have no fp32 master copy: round optimizer updates stochastically so that
                                           # updates far below a bf16 ulp (the reference's LR = 1e-5, :29) still train in expectation
    nvls: str = "auto"                     # fused MP all-reduce: auto = sum inside the NVSwitch where the box has NVLS; off = peer pushes
    master_port: int = 29500               # train_ffns.py:124
    save_path: Optional[str] = None
    load_path: Optional[str] = None
    save_every: int = 0
    profile: bool = False
    data: str = "cpu"                      # cpu = reference mock_data semantics; device = on-device Philox
    cuda_graph: bool = False               # replay the whole step from a CUDA graph (1 GPU, SGD; utils/graph.py)
    extra: dict = field(default_factory=dict)

    def __post_init__(self) -> None:
        if self.ffn_dim is None:
            self.ffn_dim = 4 * self.model_size
        if self.activation not in ("relu", "gelu"):
            raise ValueError("activation must be relu or gelu")
        if self.optimizer not in ("sgd", "adam"):
            raise ValueError("optimizer must be sgd or adam")
        if self.mp_mode not in ("ar", "sp", "spf"):
            raise ValueError("mp_mode must be ar, sp or spf")
        if self.fsdp_gather not in ("dma", "fused"):
            raise ValueError("fsdp_gather must be dma or fused")
        if self.nvls not in ("auto", "off"):
            raise ValueError("nvls must be auto or off")
        if self.mp_input not in ("full", "sharded"):
            raise ValueError("mp_input must be full or sharded")
        parse_dtype(self.dtype)

    @property
    def tokens(self) -> int:
        return self.batch_size * self.seq_len          # train_ffns.py:379

    @property
    def torch_dtype(self) -> torch.dtype:
        return parse_dtype(self.dtype)

    def to_dict(self) -> dict:
        return asdict(self)
