"""Multi-process plumbing tests on CPU/gloo (world_size 2): the reference's method-0 check (DDP == FSDP,
train_ffns.py:386-391) plus the ones it never makes (MP == single GPU; every strategy == a closed-form oracle)."""
import pytest
import torch

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.parallel import api
from distributed_llm_code_samples_b200.utils import data

from ref_oracle import load_reference

REF = load_reference()
D, FF, L, T = 16, 64, 3, 12


def _cfg(**kw):
    return TrainConfig(model_size=D, ffn_dim=FF, layers=L, device="cpu", backend="gloo", world_size=2, impl="torch",
                       lr=0.05, **kw)


def _init(bias=False):
    g = torch.Generator().manual_seed(17)
    return ffn.init_stack(L, D, FF, g, bias=bias), torch.tensor([5, 9, 13, 2])


def _oracle_dp(lp, seeds, world, lr, act="relu", residual=False):
    """Global-batch SGD with SUMMED gradients: what DDP/FSDP must equal (train_ffns.py:165, :255)."""
    ops = __import__("distributed_llm_code_samples_b200.ops", fromlist=["get_ops"]).get_ops("torch", "cpu")
    lp = [[p.clone() for p in l] for l in lp]
    opts = ffn.BlockOpts(activation=act, residual=residual)
    for step_seeds in seeds.reshape(-1, world):
        tot = None
        for s in step_seeds.tolist():
            x, dy = data.mock_batch(s, T, D)
            _, acts = ffn.tlayers_ffn_fwd(ops, lp, x, opts)
            g, _ = ffn.tlayers_ffn_bkwd(ops, dy, lp, acts, opts)
            tot = g if tot is None else [tuple(a + b for a, b in zip(ga, gb)) for ga, gb in zip(tot, g)]
        for l, gl in zip(lp, tot):
            for p, gg in zip(l, gl):
                p.add_(gg, alpha=-lr)
    return lp


def _close(a, b, atol=1e-5):
    for la, lb in zip(a, b):
        for p, q in zip(la, lb):
            assert torch.allclose(p, q, atol=atol), (p - q).abs().max()


@pytest.mark.timeout(300)
def test_ddp_equals_fsdp_equals_oracle():
    lp, seeds = _init()
    ddp = api.train_ddp(lp, seeds, T, D, _cfg())
    fsdp = api.train_fsdp(lp, seeds, T, D, _cfg())
    _close(ddp, fsdp)                                 # the reference's soft assertion
    _close(ddp, _oracle_dp(lp, seeds, 2, 0.05))
    assert api.last_stats()["world"] == 2 and api.last_stats()["tokens_per_s"] > 0


@pytest.mark.timeout(300)
@pytest.mark.parametrize("mp_mode", ["ar", "sp", "spf"])
def test_mp_equals_single(mp_mode):
    lp, seeds = _init()
    single = api.train_1gpu(lp, seeds, T, D, _cfg())
    mp = api.train_tp(lp, seeds, T, D, _cfg(mp_mode=mp_mode))
    _close(single, mp)


@pytest.mark.timeout(300)
def test_extensions_gelu_residual_bias_adam():
    lp, seeds = _init(bias=True)
    kw = dict(activation="gelu", residual=True, bias=True)
    ddp = api.train_ddp(lp, seeds, T, D, _cfg(**kw))
    fsdp = api.train_fsdp(lp, seeds, T, D, _cfg(**kw))
    _close(ddp, fsdp)
    single = api.train_1gpu(lp, seeds, T, D, _cfg(**kw))
    mp = api.train_tp(lp, seeds, T, D, _cfg(**kw))
    _close(single, mp, atol=2e-5)
    a1 = api.train_ddp(lp, seeds, T, D, _cfg(optimizer="adam", bias=True))
    a2 = api.train_fsdp(lp, seeds, T, D, _cfg(optimizer="adam", bias=True))
    _close(a1, a2, atol=1e-4)


@pytest.mark.timeout(300)
@pytest.mark.skipif(REF is None, reason="reference source not present")
def test_ddp_matches_reference_functions():
    """Simulate the reference's train_process_ddp arithmetic with ITS functions on CPU (no NCCL needed)."""
    lp, seeds = _init()
    ref = [[p.clone() for p in l] for l in lp]
    for step_seeds in seeds.reshape(-1, 2):
        grads = []
        for s in step_seeds:
            (x, dy), = list(REF.mock_data(s.reshape(1), T, D))
            _, acts = REF.tlayers_ffn_fwd(ref, x)
            g, _ = REF.tlayers_ffn_bkwd(dy, ref, acts)
            grads.append(g)
        for i, l in enumerate(ref):
            for j, p in enumerate(l):
                p.add_(-0.05 * (grads[0][i][j] + grads[1][i][j]))      # all_reduce SUM then param.add_ (:165,:172)
    ours = api.train_ddp(lp, seeds, T, D, _cfg())
    _close(ours, ref)


def test_divisibility_asserts():
    lp, _ = _init()
    with pytest.raises(AssertionError):
        api.train_ddp(lp, torch.tensor([1, 2, 3]), T, D, _cfg())       # train_ffns.py:175


@pytest.mark.timeout(300)
def test_failed_rank_fails_the_launch():
    from distributed_llm_code_samples_b200.parallel.launcher import launch
    with pytest.raises(RuntimeError, match="ranks failed"):
        launch(2, _boom, (), backend="gloo", device="cpu")


def _boom(rank, world, dev):
    if rank == 1:
        raise ValueError("injected failure")


@pytest.mark.timeout(600)
@pytest.mark.parametrize("first,second", [("ddp", "fsdp"), ("fsdp", "mp"), ("mp", "ddp")])
def test_adam_state_survives_a_cross_strategy_resume(first, second, tmp_path):
    """Checkpoints carry the optimizer: one Adam step under `first`, save, resume under `second` for the rest
    == the uninterrupted run.  The moments travel in the parameters' full layout (parallel/opt_state.py), so the resuming
    strategy re-shards them like the weights.  Without the state the resumed run restarts bias correction and diverges."""
    from distributed_llm_code_samples_b200.utils.checkpoint import load_checkpoint, save_checkpoint
    lp, seeds = _init(bias=True)
    fns = {"ddp": api.train_ddp, "fsdp": api.train_fsdp, "mp": api.train_tp}
    per_step = lambda m: 1 if m == "mp" else 2                 # seeds consumed per optimizer step at world 2
    cfg = _cfg(optimizer="adam", bias=True, activation="gelu", residual=True)
    cfg.lr = 1e-2
    # uninterrupted: the data of step 1 under `first`, then of step 2 under `second`, in one optimizer history
    s1, s2 = seeds[:per_step(first)], seeds[2:2 + per_step(second)]
    a = fns[first](lp, s1, T, D, cfg)
    st = api.last_optimizer_state()
    assert st["kind"] == "adam" and st["t"] == 1 and all(m.abs().sum() > 0 for l in st["m"] for m in l)
    path = str(tmp_path / "ck.pt")
    save_checkpoint(path, a, step=1, args=cfg.to_dict(), seeds=seeds, optimizer=st)
    ck = load_checkpoint(path)
    resumed = fns[second](ck["layers_params"], s2, T, D, cfg, ck["optimizer"])
    assert api.last_optimizer_state()["t"] == 2
    # reference: both steps under single-device Adam on the same global batches (sum of the per-seed gradients)
    cold = fns[second](ck["layers_params"], s2, T, D, cfg)              # same weights, moments restarted: must differ
    one = api.train_1gpu(lp, s1[:1], T, D, cfg) if per_step(first) == 1 else None
    if one is not None and per_step(second) == 1:
        straight = api.train_1gpu(one, s2, T, D, cfg, api.last_optimizer_state())
        for la, lb in zip(straight, resumed):
            for p, q in zip(la, lb):
                assert torch.allclose(p, q, atol=5e-5), (p - q).abs().max()
    diff = max((p - q).abs().max().item() for la, lb in zip(cold, resumed) for p, q in zip(la, lb))
    assert diff > 1e-4, "resuming with the saved moments must differ from restarting them"
    for l in resumed:
        for p in l:
            assert torch.isfinite(p).all()


@pytest.mark.timeout(600)
def test_adam_ddp_then_fsdp_equals_uninterrupted_ddp():
    """Two Adam steps under DDP == one under DDP + one under FSDP resumed from the exported state (exact same math)."""
    lp, seeds = _init(bias=True)
    cfg = _cfg(optimizer="adam", bias=True)
    cfg.lr = 1e-2
    straight = api.train_ddp(lp, seeds, T, D, cfg)
    half = api.train_ddp(lp, seeds[:2], T, D, cfg)
    resumed = api.train_fsdp(half, seeds[2:], T, D, cfg, api.last_optimizer_state())
    for la, lb in zip(straight, resumed):
        for p, q in zip(la, lb):
            assert torch.allclose(p, q, atol=2e-5), (p - q).abs().max()


@pytest.mark.timeout(600)
def test_periodic_checkpoints_carry_the_optimizer(tmp_path):
    """--save_every: the intermediate checkpoint holds Adam's moments too; resuming from it (under another strategy) lands on
    the same weights as the uninterrupted run."""
    from distributed_llm_code_samples_b200.utils.checkpoint import load_checkpoint
    lp, seeds = _init()
    cfg = _cfg(optimizer="adam")
    cfg.lr = 1e-2
    cfg.save_path, cfg.save_every = str(tmp_path / "ck"), 1
    straight = api.train_ddp(lp, seeds, T, D, cfg)                       # 2 optimizer steps at world 2, checkpoint after each
    ck = load_checkpoint(str(tmp_path / "ck.step2"))                     # after the first optimizer step (2 seeds consumed)
    assert ck["step"] == 2 and ck["optimizer"]["kind"] == "adam" and ck["optimizer"]["t"] == 1
    cfg2 = _cfg(optimizer="adam")
    cfg2.lr = 1e-2
    resumed = api.train_fsdp(ck["layers_params"], seeds[2:], T, D, cfg2, ck["optimizer"])
    for la, lb in zip(straight, resumed):
        for p, q in zip(la, lb):
            assert torch.allclose(p, q, atol=2e-5), (p - q).abs().max()
