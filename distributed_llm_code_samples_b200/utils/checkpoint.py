"""Checkpoint / resume in a plain torch ``state_dict`` layout (the reference never saves anything; layout from
SURVEY.md 5.4).  The file always holds the FULL, unsharded model

    {"model": {"layers.{i}.ffn1.weight": [F, D], "layers.{i}.ffn2.weight": [D, F], ...},
     "optimizer": {...}, "step": int, "args": {...}, "seeds": LongTensor}

so a checkpoint written under one strategy resumes under any other: resharding is just the chunk rules of
``train_ffns.py:266`` (FSDP, dim 0) and ``:317-319`` (MP, dims 0/1).  ``model`` loads into
``models.ffn.FFNStack`` (an ``nn.Module`` of ``nn.Linear``) with ``load_state_dict``.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence

import torch

from ..models.ffn import Params, params_to_state_dict, state_dict_to_params


def save_checkpoint(path: str, layers_params: Sequence[Params], step: int, args: Optional[dict] = None,
                    seeds: Optional[torch.Tensor] = None, optimizer: Optional[dict] = None) -> None:
    sd = {k: v.detach().to("cpu") for k, v in params_to_state_dict(layers_params).items()}
    payload = {"model": sd, "optimizer": optimizer or {}, "step": int(step), "args": args or {}, "seeds": seeds}
    tmp = path + ".tmp"
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(payload, tmp)
    os.replace(tmp, path)           # atomic: a crash never leaves a half-written checkpoint


def load_checkpoint(path: str) -> dict:
    ck = torch.load(path, map_location="cpu", weights_only=False)
    ck["layers_params"] = state_dict_to_params(ck["model"])
    return ck


def remaining_seeds(seeds: torch.Tensor, step: int, consumed_per_step: int = 1) -> torch.Tensor:
    """Resume continues the seed vector where the run stopped (``train_ffns.py:360``)."""
    return seeds[step * consumed_per_step:]
