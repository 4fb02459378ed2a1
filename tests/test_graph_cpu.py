"""utils/graph.py refuses what it cannot capture (the capture itself needs a GPU: tests/test_train_gpu.py)."""
import pytest
import torch

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import get_ops
from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
from distributed_llm_code_samples_b200.utils.graph import GraphedStep


def test_graphed_step_needs_cuda_tensors():
    cfg = TrainConfig(model_size=8, layers=1)
    tr = SingleTrainer(get_ops("torch", "cpu"), cfg, ffn.init_stack(1, 8, 32, torch.Generator().manual_seed(0)))
    with pytest.raises(ValueError, match="CUDA"):
        GraphedStep(tr, torch.zeros(4, 8), torch.zeros(4, 8))


def test_cuda_graph_flag_is_ignored_on_cpu():
    """--cuda_graph on a CPU run falls through to the eager step (same result as without the flag)."""
    from distributed_llm_code_samples_b200.parallel import api
    g = torch.Generator().manual_seed(1)
    lp = ffn.init_stack(2, 8, 32, g)
    seeds = torch.tensor([3, 4])
    a = api.train_1gpu(lp, seeds, 16, 8, TrainConfig(model_size=8, device="cpu", cuda_graph=True))
    b = api.train_1gpu(lp, seeds, 16, 8, TrainConfig(model_size=8, device="cpu"))
    for la, lb in zip(a, b):
        for p, q in zip(la, lb):
            assert torch.equal(p, q)
