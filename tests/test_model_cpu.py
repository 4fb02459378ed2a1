"""CPU oracle tests: our ops/model vs the reference's own pure functions and vs autograd."""
import pytest
import torch

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import get_ops
from distributed_llm_code_samples_b200.utils import data

from ref_oracle import load_reference

REF = load_reference()
needs_ref = pytest.mark.skipif(REF is None, reason="reference source not present")
OPS = get_ops("torch", "cpu")


def _stack(L=3, D=16, F=64, seed=5, bias=False):
    g = torch.Generator().manual_seed(seed)
    return ffn.init_stack(L, D, F, g, bias=bias)


@needs_ref
def test_init_matches_reference():
    g1, g2 = torch.Generator().manual_seed(3), torch.Generator().manual_seed(3)
    ours = ffn.init_tlayer_ffn(8, 32, g1)
    ref = REF.init_tlayer_ffn(8, 32, g2)
    assert [tuple(p.shape) for p in ours] == [(32, 8), (8, 32)]
    for a, b in zip(ours, ref):
        assert torch.equal(a, b)


@needs_ref
def test_mock_data_bit_equal():
    seeds = torch.tensor([11, 7, 11])
    ours = list(data.mock_data(seeds, 12, 8))
    ref = list(REF.mock_data(seeds, 12, 8))
    for (x, d), (xr, dr) in zip(ours, ref):
        assert torch.equal(x, xr) and torch.equal(d, dr)
    assert torch.equal(ours[0][0], ours[2][0])          # same seed => same batch


def test_seed_split_is_strided():
    seeds = torch.arange(4)
    parts = data.split_seeds(seeds, 2)
    assert [p.tolist() for p in parts] == [[0, 2], [1, 3]]      # train_ffns.py:182
    with pytest.raises(ValueError):
        data.split_seeds(torch.arange(3), 2)


@needs_ref
@pytest.mark.parametrize("recompute", [False, True])
def test_block_fwd_bwd_matches_reference(recompute):
    p = _stack(1)[0]
    x = torch.randn(10, 16)
    dy = torch.randn(10, 16)
    opts = ffn.BlockOpts(recompute=recompute)
    y, saved = ffn.tlayer_ffn_fwd(OPS, p, x, opts)
    assert torch.allclose(y, REF.tlayer_ffn_fwd(p, x), atol=1e-6)
    dx, (dW1, dW2) = ffn.tlayer_ffn_bkwd(OPS, dy, p, saved, opts)
    rdx, (rdW1, rdW2) = REF.tlayer_ffn_bkwd(dy.clone(), p, x)
    assert torch.allclose(dx, rdx, atol=1e-5) and torch.allclose(dW1, rdW1, atol=1e-5) and torch.allclose(dW2, rdW2, atol=1e-5)


@needs_ref
def test_stack_matches_reference():
    lp = _stack(3)
    x, dy = torch.randn(6, 16), torch.randn(6, 16)
    y, acts = ffn.tlayers_ffn_fwd(OPS, lp, x)
    ry, racts = REF.tlayers_ffn_fwd(lp, x)
    assert torch.allclose(y, ry, atol=1e-6)
    calls = []
    grads, handles = ffn.tlayers_ffn_bkwd(OPS, dy, lp, acts, after_comms_hook=lambda i, g: calls.append(i) or i)
    rgrads, _ = REF.tlayers_ffn_bkwd(dy.clone(), lp, racts)
    assert calls == [2, 1, 0] and handles == [0, 1, 2]           # reverse walk, results in layer order (:87,:94)
    for g, rg in zip(grads, rgrads):
        assert torch.allclose(g[0], rg[0], atol=1e-5) and torch.allclose(g[1], rg[1], atol=1e-5)


@pytest.mark.parametrize("act", ["relu", "gelu"])
@pytest.mark.parametrize("residual", [False, True])
@pytest.mark.parametrize("bias", [False, True])
def test_manual_backward_matches_autograd(act, residual, bias):
    lp = _stack(2, bias=bias)
    if bias:
        for p in lp:
            p[2].normal_(); p[3].normal_()
    model = ffn.FFNStack.from_params(lp, activation=act, residual=residual).double()
    x = torch.randn(5, 16, dtype=torch.double, requires_grad=True)
    dy = torch.randn(5, 16, dtype=torch.double)
    model(x).backward(dy)
    lp64 = [[p.double() for p in l] for l in lp]
    opts = ffn.BlockOpts(activation=act, residual=residual)
    y, acts = ffn.tlayers_ffn_fwd(OPS, lp64, x.detach(), opts)
    grads, _ = ffn.tlayers_ffn_bkwd(OPS, dy, lp64, acts, opts, skip_first_dx=False)
    for i, blk in enumerate(model.layers):
        assert torch.allclose(grads[i][0], blk.ffn1.weight.grad, atol=1e-8)
        assert torch.allclose(grads[i][1], blk.ffn2.weight.grad, atol=1e-8)
        if bias:
            assert torch.allclose(grads[i][2].double(), blk.ffn1.bias.grad, atol=1e-5)
            assert torch.allclose(grads[i][3].double(), blk.ffn2.bias.grad, atol=1e-5)


def test_fused_update_order_equals_separate_update():
    """Updating W inside the dW step (input-gradient first) gives the same result as grads-then-update."""
    lp = _stack(2)
    lp2 = [[p.clone() for p in l] for l in lp]
    x, dy = torch.randn(6, 16), torch.randn(6, 16)
    _, acts = ffn.tlayers_ffn_fwd(OPS, lp, x)
    grads, _ = ffn.tlayers_ffn_bkwd(OPS, dy, lp, acts)
    for l, g in zip(lp, grads):
        for p, gg in zip(l, g):
            p.add_(gg, alpha=-0.1)
    _, acts2 = ffn.tlayers_ffn_fwd(OPS, lp2, x)
    d = dy
    for i in reversed(range(2)):
        d, g = ffn.tlayer_ffn_bkwd(OPS, d, lp2[i], acts2[i], need_dx=i > 0, update_lr=0.1)
        assert g is None
    for l, l2 in zip(lp, lp2):
        for p, q in zip(l, l2):
            assert torch.allclose(p, q, atol=1e-6)


def test_state_dict_layout_roundtrip(tmp_path):
    from distributed_llm_code_samples_b200.utils.checkpoint import load_checkpoint, save_checkpoint
    lp = _stack(2, bias=True)
    sd = ffn.params_to_state_dict(lp)
    assert set(sd) == {f"layers.{i}.ffn{k}.{w}" for i in range(2) for k in (1, 2) for w in ("weight", "bias")}
    assert sd["layers.0.ffn1.weight"].shape == (64, 16) and sd["layers.0.ffn2.weight"].shape == (16, 64)
    path = str(tmp_path / "ck.pt")
    save_checkpoint(path, lp, step=3, args=TrainConfig(model_size=16).to_dict(), seeds=torch.arange(5))
    ck = load_checkpoint(path)
    assert ck["step"] == 3 and torch.equal(ck["seeds"], torch.arange(5))
    for a, b in zip(lp, ck["layers_params"]):
        for p, q in zip(a, b):
            assert torch.equal(p, q)
    m = ffn.FFNStack(2, 16, 64, bias=True)
    m.load_state_dict(ck["model"])          # loads into a plain nn.Module of nn.Linear


@needs_ref
def test_single_trainer_matches_reference_train_1gpu_math():
    """``train_1gpu`` semantics (train_ffns.py:101-116) re-derived on CPU from the reference's own fwd/bwd."""
    from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
    lp = _stack(2)
    ref = [[p.clone() for p in l] for l in lp]
    seeds = torch.tensor([4, 9, 2])
    for x, dy in REF.mock_data(seeds, 8, 16):
        _, acts = REF.tlayers_ffn_fwd(ref, x)
        g, _ = REF.tlayers_ffn_bkwd(dy, ref, acts)
        ref = [[p - REF.LR * gg for p, gg in zip(l, gl)] for l, gl in zip(ref, g)]
    tr = SingleTrainer(OPS, TrainConfig(model_size=16), [[p.clone() for p in l] for l in lp])
    tr.run(data.mock_data(seeds, 8, 16))
    for a, b in zip(tr.params, ref):
        for p, q in zip(a, b):
            assert torch.allclose(p, q, atol=1e-7)
