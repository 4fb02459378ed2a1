"""The ``mp_mode="spf"`` schedule (parallel/fused.py::spf_step) on CPU: its three primitives - all-gather+GEMM,
GEMM+reduce-scatter, final all-gather - are replaced by gloo stand-ins with the SAME contracts (in-place full-size
buffers, own token rows only final), so buffer/row/layout mistakes in the schedule show up without a GPU.  The result
must equal single-device training (as the reference's TP does, train_ffns.py:296-312)."""
import pytest
import torch
import torch.distributed as dist

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import get_ops, native as nv
from distributed_llm_code_samples_b200.parallel import api
from distributed_llm_code_samples_b200.parallel.base import block_opts, tp_shard
from distributed_llm_code_samples_b200.parallel.fused import spf_step
from distributed_llm_code_samples_b200.parallel.launcher import launch
from distributed_llm_code_samples_b200.utils import data

D, FF, L, T = 16, 64, 3, 12


class _GlooSpf:
    """State + primitives `spf_step` expects, on CPU tensors and gloo collectives."""

    def __init__(self, rank, world, shards, cfg):
        self.rank, self.world, self.shards, self.L = rank, world, shards, len(shards)
        self.ops, self.opts, self.lr = get_ops("torch", "cpu"), block_opts(cfg), cfg.lr
        nan = float("nan")            # rows that were never gathered must not be read: poison them
        self.X = [torch.full((T, D), nan) for _ in range(self.L + 1)]
        self.DX = [torch.full((T, D), nan) for _ in range(2)]

    def update_local(self, W, dy, x):
        self.ops.grad_weight_sgd_(W, dy, x, self.lr)

    def _gather_rows(self, buf):
        Ts = buf.shape[0] // self.world
        parts = [torch.empty(Ts, buf.shape[1]) for _ in range(self.world)]
        dist.all_gather(parts, buf[self.rank * Ts:(self.rank + 1) * Ts].contiguous())
        for q, p in enumerate(parts):
            if q != self.rank:
                buf[q * Ts:(q + 1) * Ts] = p

    def update_bias(self, layer, dh, dy):
        if len(self.shards[layer]) == 4:
            self.shards[layer][2].sub_(self.lr * dh.sum(0))
            self.shards[layer][3].sub_(self.lr * dy.sum(0))

    def ag_gemm(self, buf, W, b_major, kind, aux=None, bias=None):
        self._gather_rows(buf)
        assert torch.isfinite(buf).all()
        h = buf @ (W.t() if b_major == nv.MAJOR_K else W)
        if bias is not None:
            h = h + bias
        if kind == "relu":
            a = torch.where(h <= 0, torch.zeros_like(h), h)
            return a, a
        if kind == "gelu":
            return torch.nn.functional.gelu(h), h
        if kind == "relu_bwd":
            return h.masked_fill(aux <= 0, 0.0), None
        if kind == "gelu_bwd":
            pre = aux.detach().requires_grad_(True)
            g, = torch.autograd.grad(torch.nn.functional.gelu(pre).sum(), pre)
            return h * g, None
        raise ValueError(kind)

    def rs_gemm(self, A, W, b_major, dst, residual=None, bias=None):
        full = A @ (W.t() if b_major == nv.MAJOR_K else W)
        dist.all_reduce(full)
        if bias is not None:
            full = full + bias
        Ts = dst.shape[0] // self.world
        rows = slice(self.rank * Ts, (self.rank + 1) * Ts)
        res = residual[rows].clone() if residual is not None else 0.0     # (residual may alias dst's buffer family)
        dst[rows] = full[rows] + res
        # the rows of dst that belong to other ranks keep whatever they had: poison them so a read would be caught
        for q in range(self.world):
            if q != self.rank:
                dst[q * Ts:(q + 1) * Ts] = float("nan")

    def gather_out(self, buf):
        self._gather_rows(buf)
        return buf.clone()


def _worker(rank, world, dev, lp, seeds, cfg_dict, out):
    cfg = TrainConfig(**cfg_dict)
    shards = [[tp_shard(p, j, rank, world).clone() for j, p in enumerate(l)] for l in lp]
    t = _GlooSpf(rank, world, shards, cfg)
    for x, dy in data.mock_data(seeds, T, D, torch.float32, dev):
        y = spf_step(t, x, dy)
        assert torch.isfinite(y).all()
    for i, l in enumerate(shards):
        out[i][0].chunk(world, dim=0)[rank].copy_(l[0])
        out[i][1].chunk(world, dim=1)[rank].copy_(l[1])
        if len(l) == 4:
            out[i][2].chunk(world, dim=0)[rank].copy_(l[2])
            if rank == 0:
                out[i][3].copy_(l[3])


@pytest.mark.timeout(300)
@pytest.mark.parametrize("act,recompute,residual,bias", [("relu", False, False, False), ("gelu", False, True, False),
                                                         ("relu", True, True, False), ("gelu", False, True, True)])
def test_spf_schedule_equals_single_device(act, recompute, residual, bias):
    g = torch.Generator().manual_seed(17)
    lp = ffn.init_stack(L, D, FF, g, bias=bias)
    if bias:
        for l in lp:                  # non-zero biases so that a dropped or doubled bias shows
            l[2].normal_(0, 0.1, generator=g)
            l[3].normal_(0, 0.1, generator=g)
    seeds = torch.tensor([5, 9, 13])
    cfg = TrainConfig(model_size=D, ffn_dim=FF, layers=L, device="cpu", backend="gloo", world_size=2, impl="torch", lr=0.05,
                      activation=act, recompute=recompute, residual=residual, mp_mode="spf")
    shared = [[p.clone().share_memory_() for p in l] for l in lp]
    out = [[torch.empty_like(p).share_memory_() for p in l] for l in lp]
    launch(2, _worker, (shared, seeds, cfg.to_dict(), out), backend="gloo", device="cpu")
    single = api.train_1gpu(lp, seeds, T, D, cfg)
    for la, lb in zip(single, out):
        for p, q in zip(la, lb):
            assert torch.allclose(p, q, atol=2e-5), (p - q).abs().max()
