"""distributed_llm_code_samples_b200 - a Blackwell-native FFN-stack trainer (1 GPU / DDP / FSDP / MP).

Capabilities follow martin-kukla/distributed-llm-code-samples (``train_ffns.py``); the design does not: the hot
ops are hand-written sm_100a kernels (``csrc/``, C ABI in ``include/b200ffn.h``) and the collectives are fused
into the GEMMs over NVLink peer memory.  See DESIGN.md.
"""
from .config import LR, DLOSS_DX_COEF, INIT_SCALE, TrainConfig  # noqa: F401

__version__ = "0.1.0"
