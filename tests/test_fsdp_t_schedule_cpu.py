"""The transposed-activation FSDP schedule (parallel/fused_fsdp_t.py::fsdp_t_step) on CPU: its four primitives are
replaced by gloo / torch stand-ins with the same contracts - in particular the generic GEMM ``D = sum_k A(m,k) B(n,k)``
with K- or MN-major operands exactly as the native kernel defines them - so every operand layout, shape and epilogue
choice of the schedule is checked against single-device training (train_ffns.py:101-116)."""
import pytest
import torch
import torch.distributed as dist

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import native as nv
from distributed_llm_code_samples_b200.parallel import api
from distributed_llm_code_samples_b200.parallel.base import block_opts, fsdp_shard
from distributed_llm_code_samples_b200.parallel.fused_fsdp_t import fsdp_t_step
from distributed_llm_code_samples_b200.parallel.launcher import launch
from distributed_llm_code_samples_b200.utils import data

D, FF, L, T = 16, 64, 3, 12


def _mat(X, major, rows_are):
    """Operand as a [rows, K] matrix: K-major storage is already that; MN-major storage is [K, rows]."""
    return X if major == nv.MAJOR_K else X.t()


def _gemm(A, a_major, B, b_major, M, N, K, epi, aux=None):
    Am, Bm = _mat(A, a_major, M), _mat(B, b_major, N)
    assert Am.shape == (M, K) and Bm.shape == (N, K), (Am.shape, Bm.shape, M, N, K)
    acc = Am @ Bm.t()
    if epi == nv.EPI_STORE:
        return acc, None
    if epi == nv.EPI_ADD:
        return acc + aux, None
    if epi in (nv.EPI_RELU, nv.EPI_RELU_MASK):
        out = torch.where(acc <= 0, torch.zeros_like(acc), acc)
        return out, None                       # (no packed mask on CPU: the schedule then uses the output itself)
    if epi == nv.EPI_GELU:
        return torch.nn.functional.gelu(acc), acc
    if epi in (nv.EPI_RELU_BWD, nv.EPI_RELU_BWD_MASK):
        return acc.masked_fill(aux <= 0, 0.0), None
    if epi == nv.EPI_GELU_BWD:
        pre = aux.detach().requires_grad_(True)
        g, = torch.autograd.grad(torch.nn.functional.gelu(pre).sum(), pre)
        return acc * g, None
    raise ValueError(epi)


class _GlooFsdpT:
    def __init__(self, rank, world, shards, cfg):
        self.rank, self.world, self.shards, self.L = rank, world, shards, len(shards)
        self.opts, self.lr = block_opts(cfg), cfg.lr

    def _full(self, l, k):
        parts = [torch.empty_like(self.shards[l][k]) for _ in range(self.world)]
        dist.all_gather(parts, self.shards[l][k].contiguous())
        return torch.cat(parts, dim=0)

    def gather_w_gemm(self, l, k, B, N, epi, aux):
        W = self._full(l, k)
        return _gemm(W, nv.MAJOR_K, B, nv.MAJOR_MN, W.shape[0], N, W.shape[1], epi, aux)

    def bwd_w_gemm(self, l, k, B, N, epi, aux):
        W = self._full(l, k)                   # storage [K, M], read MN-major
        return _gemm(W, nv.MAJOR_MN, B, nv.MAJOR_MN, W.shape[1], N, W.shape[0], epi, aux)[0]

    def transpose(self, x):
        return x.t().contiguous()

    def update(self, l, k, A, B):
        G, _ = _gemm(A, nv.MAJOR_K, B, nv.MAJOR_K, A.shape[0], B.shape[0], A.shape[1], nv.EPI_STORE)
        dist.all_reduce(G)
        rows = G.shape[0] // self.world
        assert self.shards[l][k].shape == (rows, G.shape[1])
        self.shards[l][k].add_(G[self.rank * rows:(self.rank + 1) * rows], alpha=-self.lr)


def _worker(rank, world, dev, lp, seeds, cfg_dict, out):
    cfg = TrainConfig(**cfg_dict)
    shards = [[fsdp_shard(p, rank, world).clone() for p in l] for l in lp]
    t = _GlooFsdpT(rank, world, shards, cfg)
    mine = data.split_seeds(seeds, world)[rank]
    for x, dy in data.mock_data(mine, T, D, torch.float32, dev):
        y = fsdp_t_step(t, x, dy)
        assert y.shape == (T, D) and torch.isfinite(y).all()
    for i, l in enumerate(shards):
        for j, p in enumerate(l):
            out[i][j].chunk(world, dim=0)[rank].copy_(p)


@pytest.mark.timeout(300)
@pytest.mark.parametrize("act,residual", [("relu", False), ("gelu", True)])
def test_transposed_fsdp_schedule_equals_ddp(act, residual):
    g = torch.Generator().manual_seed(17)
    lp = ffn.init_stack(L, D, FF, g)
    seeds = torch.tensor([5, 9, 13, 2])
    cfg = TrainConfig(model_size=D, ffn_dim=FF, layers=L, device="cpu", backend="gloo", world_size=2, impl="torch", lr=0.05,
                      activation=act, residual=residual)
    shared = [[p.clone().share_memory_() for p in l] for l in lp]
    out = [[torch.empty_like(p).share_memory_() for p in l] for l in lp]
    launch(2, _worker, (shared, seeds, cfg.to_dict(), out), backend="gloo", device="cpu")
    ref = api.train_fsdp(lp, seeds, T, D, cfg)          # the plain FSDP trainer on gloo (itself checked against DDP and an oracle)
    for la, lb in zip(ref, out):
        for p, q in zip(la, lb):
            assert torch.allclose(p, q, atol=2e-5), (p - q).abs().max()
