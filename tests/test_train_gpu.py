"""End-to-end on one B200: the native single-device trainer vs the PyTorch-ops trainer (fp32 oracle of the same
math on the same bf16 inputs), smoke(), and the Engine's end-to-end loop."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
# written after round 1's GPU budget was spent: opt-in until it has run on hardware once
queued = pytest.mark.skipif(os.environ.get("B200_QUEUED_TESTS") != "1", reason="queued for the next GPU round (set B200_QUEUED_TESTS=1)")


def _setup(dev, L=3, D=256, FF=1024, T=512, seed=3):
    from distributed_llm_code_samples_b200.models import ffn
    g = torch.Generator().manual_seed(seed)
    lp = ffn.init_stack(L, D, FF, g)
    x = torch.randn(T, D, generator=g)
    dy = 0.1 * torch.randn(T, D, generator=g)
    return lp, x, dy


@pytest.mark.parametrize("recompute", [False, True])
@pytest.mark.parametrize("act", ["relu", "gelu"])
def test_native_step_matches_fp32_oracle(cuda_device, recompute, act):
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.ops import get_ops
    from distributed_llm_code_samples_b200.parallel.base import place
    from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
    lp, x, dy = _setup(cuda_device)
    lr = 0.5      # large enough for the update to register in bf16
    cfg16 = TrainConfig(model_size=256, ffn_dim=1024, dtype="bf16", lr=lr, recompute=recompute, activation=act)
    cfg32 = TrainConfig(model_size=256, ffn_dim=1024, dtype="fp32", lr=lr, recompute=recompute, activation=act)
    lp16 = [[p.to(torch.bfloat16) for p in l] for l in lp]
    nat = SingleTrainer(get_ops("native", cuda_device), cfg16, place(lp16, cuda_device, torch.bfloat16))
    ora = SingleTrainer(get_ops("torch", cuda_device), cfg32, place(lp16, cuda_device, torch.float32))
    xb, dyb = x.to(torch.bfloat16).to(cuda_device), dy.to(torch.bfloat16).to(cuda_device)
    y = nat.step(xb, dyb)
    yo = ora.step(xb.float(), dyb.float())
    torch.cuda.synchronize()
    assert nat.ops.fallbacks == 0
    assert (y.float() - yo).abs().max().item() <= 2e-2 * max(1.0, yo.abs().max().item())
    for a, b, c in zip(nat.params, ora.params, lp16):
        for p, q, p0 in zip(a, b, c):
            upd_ref = q - p0.float().to(cuda_device)
            upd = p.float() - p0.float().to(cuda_device)
            # the UPDATE must match (relative to its own scale), not just the weights.  Intermediates (a, dh, dx) and the
            # weights themselves are bf16 in the native path and fp32 in the oracle, hence a norm-based bound.
            rel = ((upd - upd_ref).norm() / upd_ref.norm()).item()
            assert rel <= 0.10, f"relative update error {rel}"
            assert (upd - upd_ref).abs().max().item() <= 0.15 * upd_ref.abs().max().item() + 2e-3


@pytest.mark.parametrize("optimizer", ["sgd", "adam"])
def test_north_star_block_native_no_fallbacks(cuda_device, optimizer):
    """Linear -> GELU -> Linear + residual with biases (BASELINE.json's block; the reference has neither bias nor residual,
    train_ffns.py:35,:54-58): native path vs the fp32 oracle, and NOT ONE op may fall back to PyTorch - bias add in the GEMM
    epilogues (fp32 resident biases), bias gradient = native column sum, fp32 bias update = native kernel."""
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.models import ffn
    from distributed_llm_code_samples_b200.ops import get_ops
    from distributed_llm_code_samples_b200.parallel.base import place
    from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
    g = torch.Generator().manual_seed(5)
    lp = ffn.init_stack(2, 256, 1024, g, bias=True)
    for l in lp:
        l[2].normal_(0, 0.05, generator=g)
        l[3].normal_(0, 0.05, generator=g)
    x, dy = torch.randn(512, 256, generator=g), 0.1 * torch.randn(512, 256, generator=g)
    lr = 0.05 if optimizer == "sgd" else 1e-2
    kw = dict(model_size=256, ffn_dim=1024, lr=lr, activation="gelu", residual=True, bias=True, optimizer=optimizer)
    lp16 = [[p.to(torch.bfloat16) if j < 2 else p for j, p in enumerate(l)] for l in lp]
    nat = SingleTrainer(get_ops("native", cuda_device), TrainConfig(dtype="bf16", **kw), place(lp16, cuda_device, torch.bfloat16))
    ora = SingleTrainer(get_ops("torch", cuda_device), TrainConfig(dtype="fp32", **kw), place(lp16, cuda_device, torch.float32))
    assert nat.params[0][2].dtype == torch.float32, "biases stay fp32 on the bf16 path"
    xb, dyb = x.to(torch.bfloat16).to(cuda_device), dy.to(torch.bfloat16).to(cuda_device)
    for _ in range(2):
        y = nat.step(xb, dyb)
        yo = ora.step(xb.float(), dyb.float())
    torch.cuda.synchronize()
    assert nat.ops.fallbacks == 0, "an op fell back to PyTorch"
    assert (y.float() - yo).abs().max().item() <= 3e-2 * max(1.0, yo.abs().max().item())
    for a, b, c in zip(nat.params, ora.params, lp16):
        for j, (p, q, p0) in enumerate(zip(a, b, c)):
            upd_ref = q.float() - p0.float().to(cuda_device)
            upd = p.float() - p0.float().to(cuda_device)
            rel = ((upd - upd_ref).norm() / upd_ref.norm()).item()
            assert rel <= (0.12 if j < 2 else 0.05), (j, rel)      # biases are fp32 on both sides


@pytest.mark.parametrize("act,residual", [("relu", False), ("gelu", True)])
def test_fp32_parity_mode_runs_on_native_kernels(cuda_device, act, residual):
    """--dtype fp32 (the reference's dtype, train_ffns.py:36,:149, and this CLI's default) must NOT fall back to cuBLAS/ATen:
    every GEMM runs on tcgen05 through split bf16 operands (csrc/gemm_f32.cu), and the result matches the PyTorch fp32 path
    to fp32 accuracy - tight enough for the reference's own DDP==FSDP allclose (rtol 1e-5) to be meaningful."""
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.ops import get_ops
    from distributed_llm_code_samples_b200.parallel.base import place
    from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
    lp, x, dy = _setup(cuda_device)
    cfg = TrainConfig(model_size=256, ffn_dim=1024, dtype="fp32", lr=0.05, activation=act, residual=residual)
    nat = SingleTrainer(get_ops("native", cuda_device), cfg, place(lp, cuda_device, torch.float32))
    ora = SingleTrainer(get_ops("torch", cuda_device), cfg, place(lp, cuda_device, torch.float32))
    # fp64 run of the same schedule: the yardstick both fp32 implementations are measured against
    o64 = SingleTrainer(get_ops("torch", cuda_device), cfg, [[p.to(cuda_device, torch.float64) for p in l] for l in lp])
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False

    def err(t, ref):
        return ((t.double() - ref).abs().max() / ref.abs().max()).item()

    try:
        xb, dyb = x.to(cuda_device), dy.to(cuda_device)
        y, yo, y64 = nat.step(xb, dyb), ora.step(xb, dyb), o64.step(xb.double(), dyb.double())
        torch.cuda.synchronize()
        assert nat.ops.fallbacks == 0, "fp32 mode fell back to PyTorch ops"
        # Measured on B200 (tools/f32_probe.py): output within 2.6e-6 of its scale after three blocks (cuBLAS fp32: ~7e-7), every
        # weight within 1.5e-7 of its own after the step (cuBLAS: ~1e-7); a TF32 GEMM would sit near 5e-4, bf16 near 4e-3.
        assert err(y, y64) <= 6 * err(yo, y64) + 2e-6, (err(y, y64), err(yo, y64))
        for a, b, c in zip(nat.params, ora.params, o64.params):
            for p, q, r in zip(a, b, c):
                assert err(p, r) <= 6 * err(q, r) + 1e-6, (err(p, r), err(q, r))
        # A second step stays finite and close.  (Weights are not compared after it: by then a pre-activation within 1e-6 of
        # zero can take the other ReLU branch than the fp64 run - one such flip moves a whole weight row by ~5e-4 of the
        # weight scale.  That is a property of comparing ANY two fp32 runs of a ReLU net, not of these kernels.)
        y2, y264 = nat.step(xb, dyb), o64.step(xb.double(), dyb.double())
        torch.cuda.synchronize()
        assert torch.isfinite(y2).all() and err(y2, y264) < 1e-4
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def test_smoke_entry(cuda_device):
    import __graft_entry__ as ge
    ge.smoke()


def test_engine_end_to_end_loop(cuda_device):
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.engine import Engine
    cfg = TrainConfig(batch_size=2, seq_len=128, layers=2, model_size=256, ffn_dim=1024, dtype="bf16", device="cuda", impl="native")
    eng = Engine(cfg, method="ddp", rank=0, world=1, device=cuda_device)
    pool = []
    for i in range(3):
        g = torch.Generator().manual_seed(i)
        pool.append((torch.randn(cfg.tokens, 256, generator=g).to(torch.bfloat16).pin_memory(),
                     (0.1 * torch.randn(cfg.tokens, 256, generator=g)).to(torch.bfloat16).pin_memory()))
    metrics = list(eng.train(iter(pool)))
    assert len(metrics) == 3 and all(m > 0 for m in metrics)
    assert eng.h2d_bytes == 3 * 2 * cfg.tokens * 256 * 2 and eng.d2h_bytes == 12


def test_cli_single_gpu_bf16(cuda_device, capsys):
    import train_ffns
    rc = train_ffns.main(["-s", "2", "-bs", "2", "-n", "128", "-l", "2", "-d", "256", "-m", "1", "-r", "7", "--dtype", "bf16"])
    out = capsys.readouterr().out
    assert rc == 0 and "train_1gpu takes" in out and "PARAMS:" in out and "tokens/s" in out


def test_cuda_graph_replay_matches_eager_steps(cuda_device):
    """utils/graph.py: three replays of the captured step leave exactly the weights three eager steps leave (same
    kernels, deterministic), and the warm-up steps inside the capture did not train."""
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.ops import get_ops
    from distributed_llm_code_samples_b200.parallel.base import place
    from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
    from distributed_llm_code_samples_b200.utils.graph import GraphedStep
    lp, x, dy = _setup(cuda_device)
    cfg = TrainConfig(model_size=256, ffn_dim=1024, dtype="bf16", lr=0.05)
    lp16 = [[p.to(torch.bfloat16) for p in l] for l in lp]
    eager = SingleTrainer(get_ops("native", cuda_device), cfg, place(lp16, cuda_device, torch.bfloat16))
    graphed = SingleTrainer(get_ops("native", cuda_device), cfg, place(lp16, cuda_device, torch.bfloat16))
    batches = []
    for i in range(3):
        g = torch.Generator().manual_seed(50 + i)
        batches.append((torch.randn(512, 256, generator=g).to(torch.bfloat16).to(cuda_device),
                        (0.1 * torch.randn(512, 256, generator=g)).to(torch.bfloat16).to(cuda_device)))
    gs = GraphedStep(graphed, *batches[0])
    for a, b in zip(graphed.params, lp16):
        for p, q in zip(a, b):
            assert torch.equal(p.cpu(), q), "capture must leave the weights untouched"
    for xb, db in batches:
        ye = eager.step(xb, db).clone()
        yg = gs.step(xb, db)
        assert torch.equal(ye, yg)
    torch.cuda.synchronize()
    assert gs.replays == 3
    for a, b in zip(eager.params, graphed.params):
        for p, q in zip(a, b):
            assert torch.equal(p, q)


def test_prefetcher_row_sliced_staging(cuda_device):
    """Sequence-parallel MP stages only its token shard: those rows arrive, and only their bytes are counted."""
    from distributed_llm_code_samples_b200.utils.data import PinnedPrefetcher
    batches = [(torch.randn(64, 32).to(torch.bfloat16), torch.randn(64, 32).to(torch.bfloat16)) for _ in range(3)]
    pf = PinnedPrefetcher(iter(batches), cuda_device, rows=slice(16, 32))
    for (hx, hd), (x, d) in zip(batches, pf):
        torch.cuda.synchronize()
        assert x.shape == hx.shape and torch.equal(x[16:32].cpu(), hx[16:32]) and torch.equal(d[16:32].cpu(), hd[16:32])
        assert pf.last_bytes == 2 * 16 * 32 * 2
