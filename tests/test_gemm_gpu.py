"""tcgen05 GEMM numerics vs a plain fp32 PyTorch reference (all operand layouts, all epilogues, ragged shapes)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributed_llm_code_samples_b200.ops import native as nv  # noqa: E402


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).to(dev)


def _check(out, ref, K, what):
    out = out.float()
    err = (out - ref).abs().max().item()
    tol = 2e-2 * max(1.0, ref.abs().max().item()) + 1e-3
    assert torch.isfinite(out).all(), f"{what}: non-finite output"
    assert err <= tol, f"{what}: max abs err {err} > {tol} (K={K})"


SHAPES = [(128, 256, 64), (256, 512, 256), (200, 264, 136), (384, 768, 1024), (1024, 256, 4096), (130, 8, 72)]


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_tn(cuda_device, M, N, K):
    """y = x @ W^T (train_ffns.py:42): both operands K-major."""
    x, w = _rand((M, K), cuda_device, seed=1), _rand((N, K), cuda_device, 0.1, seed=2)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    nv.gemm(x, nv.MAJOR_K, w, nv.MAJOR_K, out, M, N, K)
    _check(out, x.float() @ w.float().T, K, "TN")


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_nn(cuda_device, M, N, K):
    """dx = dy @ W (train_ffns.py:45 'bc,cd->bd'): B is MN-major."""
    dy, w = _rand((M, K), cuda_device, seed=3), _rand((K, N), cuda_device, 0.1, seed=4)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    nv.gemm(dy, nv.MAJOR_K, w, nv.MAJOR_MN, out, M, N, K)
    _check(out, dy.float() @ w.float(), K, "NN")


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_nt(cuda_device, M, N, K):
    """dW = dy^T @ x (train_ffns.py:45 'bc,bd->cd'): both operands MN-major, K = tokens."""
    if M % 8:
        pytest.skip("MN-major A needs M % 8 == 0 (TMA pitch)")
    dy, x = _rand((K, M), cuda_device, seed=5), _rand((K, N), cuda_device, 0.1, seed=6)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    nv.gemm(dy, nv.MAJOR_MN, x, nv.MAJOR_MN, out, M, N, K)
    _check(out, dy.float().T @ x.float(), K, "NT")


def test_epilogues(cuda_device):
    M, N, K = 384, 512, 320
    x, w = _rand((M, K), cuda_device, seed=7), _rand((N, K), cuda_device, 0.1, seed=8)
    aux = _rand((M, N), cuda_device, seed=9)
    bias = torch.randn(N, device=cuda_device)
    ref = x.float() @ w.float().T
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    out2 = torch.empty_like(out)

    nv.gemm(x, 0, w, 0, out, M, N, K, nv.EPI_RELU)
    _check(out, torch.relu(ref), K, "relu")
    nv.gemm(x, 0, w, 0, out, M, N, K, nv.EPI_RELU, bias=bias)
    _check(out, torch.relu(ref + bias), K, "relu+bias")
    nv.gemm(x, 0, w, 0, out, M, N, K, nv.EPI_GELU, out2=out2)
    _check(out, torch.nn.functional.gelu(ref), K, "gelu")
    _check(out2, ref, K, "gelu pre-activation")
    nv.gemm(x, 0, w, 0, out, M, N, K, nv.EPI_RELU_BWD, aux=aux)
    _check(out, ref * (aux.float() > 0), K, "relu_bwd")
    nv.gemm(x, 0, w, 0, out, M, N, K, nv.EPI_GELU_BWD, aux=aux)
    a = aux.float().requires_grad_(True)
    torch.nn.functional.gelu(a).sum().backward()
    _check(out, ref * a.grad, K, "gelu_bwd")
    nv.gemm(x, 0, w, 0, out, M, N, K, nv.EPI_ADD, aux=aux)
    _check(out, ref + aux.float(), K, "add")
    wgt = _rand((M, N), cuda_device, seed=10)
    w0 = wgt.float().clone()
    nv.gemm(x, 0, w, 0, wgt, M, N, K, nv.EPI_SGD, alpha=0.05)
    _check(wgt, w0 - 0.05 * ref, K, "sgd")
    o32 = torch.empty(M, N, dtype=torch.float32, device=cuda_device)
    nv.gemm(x, 0, w, 0, o32, M, N, K, nv.EPI_STORE_F32, alpha=0.5)
    assert (o32 - 0.5 * ref).abs().max().item() < 1e-2


@pytest.mark.parametrize("M,N,K", [(384, 512, 320), (200, 384, 136), (1024, 1280, 512), (8192, 384, 136), (16384, 5504, 64)])
def test_relu_bitmask_epilogues(cuda_device, M, N, K):
    """ReLU forward that also packs the sign mask (1 bit/element) and the backward that consumes it (:47-52)."""
    x, w = _rand((M, K), cuda_device, seed=31), _rand((N, K), cuda_device, 0.1, seed=32)
    ref = x.float() @ w.float().T
    a = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    mask = nv.relu_mask_empty(M, N, cuda_device)
    nv.gemm(x, 0, w, 0, a, M, N, K, nv.EPI_RELU_MASK, out2=mask)
    _check(a, torch.relu(ref), K, "relu+mask")
    bits = nv.unpack_relu_mask(mask, M, N)
    assert torch.equal(nv.unpack_relu_mask(nv.pack_relu_mask(bits), M, N), bits)          # the helpers mirror the kernels' layout
    assert torch.equal(bits, a > 0)
    dy, w2 = _rand((M, K), cuda_device, seed=33), _rand((K, N), cuda_device, 0.1, seed=34)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
    nv.gemm(dy, 0, w2, 1, out, M, N, K, nv.EPI_RELU_BWD_MASK, aux=mask)
    _check(out, (dy.float() @ w2.float()) * bits, K, "relu_bwd from mask")


def _tile_rel_err(out, ref):
    """max over 128x256 tiles of (max |out - ref| in the tile) / (max |ref| in the tile)."""
    M, N = ref.shape
    d = (out.float() - ref).abs().reshape(M // 128, 128, N // 256, 256).amax(dim=(1, 3))
    s = ref.abs().reshape(M // 128, 128, N // 256, 256).amax(dim=(1, 3)).clamp_min(1e-30)
    return float((d / s).max())


@pytest.mark.parametrize("layout", ["tn", "nn", "nt", "relu_mask", "relu_bwd_mask", "sgd"])
def test_gemm_benchmark_size(cuda_device, layout):
    """Numerics at the benchmark's size, 16384 x 16384 x 4096 (148 CTAs x 14 pair tiles each, several rasterisation
    groups), per 128x256 tile against torch's fp32 GEMM of the same bf16 inputs - every layout of the training step."""
    M, N, K = 16384, 16384, 4096
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=cuda_device)
        if layout in ("tn", "relu_mask"):
            a, b = _rand((M, K), cuda_device, seed=41), _rand((N, K), cuda_device, 0.1, seed=42)
            ref = a.float() @ b.float().T
            if layout == "tn":
                nv.gemm(a, 0, b, 0, out, M, N, K)
            else:
                mask = nv.relu_mask_empty(M, N, cuda_device)
                nv.gemm(a, 0, b, 0, out, M, N, K, nv.EPI_RELU_MASK, out2=mask)
                ref = torch.relu(ref)
                bits = nv.unpack_relu_mask(mask, M, N)
                assert torch.equal(bits, out > 0)
        elif layout in ("nn", "relu_bwd_mask"):
            a, b = _rand((M, K), cuda_device, seed=43), _rand((K, N), cuda_device, 0.1, seed=44)
            ref = a.float() @ b.float()
            if layout == "nn":
                nv.gemm(a, 0, b, 1, out, M, N, K)
            else:
                keep = torch.rand(M, N, device=cuda_device) > 0.5
                mask = nv.pack_relu_mask(keep)
                nv.gemm(a, 0, b, 1, out, M, N, K, nv.EPI_RELU_BWD_MASK, aux=mask)
                ref = ref * keep
        else:
            a, b = _rand((K, M), cuda_device, seed=45), _rand((K, N), cuda_device, 0.1, seed=46)
            ref = a.float().T @ b.float()
            if layout == "nt":
                nv.gemm(a, 1, b, 1, out, M, N, K)
            else:
                out.copy_(_rand((M, N), cuda_device, 0.5, seed=47))
                w0 = out.float().clone()
                nv.gemm(a, 1, b, 1, out, M, N, K, nv.EPI_SGD, alpha=0.01)
                ref = w0 - 0.01 * ref
        torch.cuda.synchronize()
        assert torch.isfinite(out.float()).all()
        assert _tile_rel_err(out, ref) <= 1e-2, (layout, _tile_rel_err(out, ref))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    assert nv.device_error()[0] == 0


def test_gemm_strided_views(cuda_device):
    """Operands that are column slices of bigger tensors (leading dimension != row length)."""
    M, N, K = 256, 256, 128
    xb, wb = _rand((M, 2 * K), cuda_device, seed=11), _rand((N, 3 * K), cuda_device, 0.1, seed=12)
    x, w = xb[:, K:], wb[:, K:2 * K]
    outb = torch.zeros(M, 2 * N, dtype=torch.bfloat16, device=cuda_device)
    nv.gemm(x, 0, w, 0, outb[:, N:], M, N, K)
    _check(outb[:, N:], x.float() @ w.float().T, K, "strided")
    assert outb[:, :N].abs().max().item() == 0


def test_gemm_rejects_bad_shapes(cuda_device):
    x = torch.zeros(16, 12, dtype=torch.bfloat16, device=cuda_device)
    w = torch.zeros(16, 12, dtype=torch.bfloat16, device=cuda_device)
    out = torch.zeros(16, 16, dtype=torch.bfloat16, device=cuda_device)
    with pytest.raises(nv.NativeError):
        nv.gemm(x, 0, w, 0, out, 16, 16, 12)   # pitch 12 elements = 24 bytes: not a TMA-legal stride


def test_elementwise(cuda_device):
    n = 1000003
    p, g = _rand((n,), cuda_device, seed=13), _rand((n,), cuda_device, seed=14)
    p0 = p.float().clone()
    nv.sgd_(p, g, 0.1)
    assert (p.float() - (p0 - 0.1 * g.float())).abs().max().item() < 2e-2
    x = _rand((n,), cuda_device, seed=15)
    assert torch.equal(nv.relu(x), torch.relu(x))
    dy = _rand((n,), cuda_device, seed=16)
    ref = dy.clone().masked_fill_(x <= 0, 0)
    assert torch.equal(nv.relu_bwd_(dy, x), ref)
    a, b = _rand((n,), cuda_device, seed=17), _rand((n,), cuda_device, seed=18)
    assert (nv.add(a, b).float() - (a.float() + b.float())).abs().max().item() < 2e-2
    s = nv.sumsq(x)
    s2 = nv.sumsq(x)
    assert torch.equal(s, s2)                                   # deterministic, scratch reusable
    assert abs(s.item() - (x.float() ** 2).sum().item()) / s.item() < 1e-4
    r = nv.randn_(torch.empty(1 << 20, dtype=torch.bfloat16, device=cuda_device), seed=1234, scale=2.0)
    assert abs(r.float().mean().item()) < 0.02 and abs(r.float().std().item() - 2.0) < 0.05
    f = torch.randn(4099, device=cuda_device)
    assert torch.equal(nv.cast_to_bf16(f), f.to(torch.bfloat16))
    assert torch.equal(nv.cast_to_f32(nv.cast_to_bf16(f)), f.to(torch.bfloat16).float())


def test_adam(cuda_device):
    n = 4096
    p, g = _rand((n,), cuda_device, seed=19), _rand((n,), cuda_device, seed=20)
    m = torch.zeros(n, device=cuda_device)
    v = torch.zeros(n, device=cuda_device)
    pr = torch.nn.Parameter(p.float().clone())
    opt = torch.optim.Adam([pr], lr=1e-2)
    for step in range(1, 4):
        pr.grad = g.float().clone()
        opt.step()
        nv.adam_(p, g, m, v, 1e-2, 0.9, 0.999, 1e-8, step)
    assert (p.float() - pr.detach()).abs().max().item() < 6e-2   # three bf16 roundings at |p|~3


@pytest.mark.parametrize("n", [4096, 1000003])
def test_adam_vectorised_matches_fp32_reference(cuda_device, n):
    """8-wide Adam (16-byte parameter / moment accesses) + scalar tail vs the same formula in fp32 torch, both gradient
    dtypes; the moments must match to fp32 round-off."""
    for gdt in (torch.bfloat16, torch.float32):
        p = _rand((n,), cuda_device, seed=21)
        g = _rand((n,), cuda_device, seed=22).to(gdt)
        m, v = torch.zeros(n, device=cuda_device), torch.zeros(n, device=cuda_device)
        pr, mr, vr = p.float().clone(), torch.zeros(n, device=cuda_device), torch.zeros(n, device=cuda_device)
        for step in (1, 2):
            nv.adam_(p, g, m, v, 1e-2, 0.9, 0.999, 1e-8, step)
            mr = 0.9 * mr + 0.1 * g.float()
            vr = 0.999 * vr + 0.001 * g.float() ** 2
            pr = (pr - 1e-2 * (mr / (1 - 0.9 ** step)) / ((vr / (1 - 0.999 ** step)).sqrt() + 1e-8)).to(torch.bfloat16).float()
            assert (m - mr).abs().max().item() < 1e-5 and (v - vr).abs().max().item() < 1e-5
            assert (p.float() - pr).abs().max().item() <= 2 ** -6      # one bf16 ulp at |p| < 4
            pr = p.float().clone()


def test_fp32_param_optimizers_and_bias_kernels(cuda_device):
    """fp32 SGD / Adam (the biases), the column-sum bias gradient (two-stage, deterministic) and the bias/residual row fill."""
    n = 5000
    p, g = torch.randn(n, device=cuda_device), torch.randn(n, device=cuda_device)
    p0 = p.clone()
    nv.sgd_f32_(p, g, 0.1)
    assert torch.allclose(p, p0 - 0.1 * g, atol=1e-6)
    m, v = torch.zeros(n, device=cuda_device), torch.zeros(n, device=cuda_device)
    pr = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([pr], lr=1e-2)
    for step in (1, 2, 3):
        pr.grad = g.clone()
        opt.step()
        nv.adam_f32_(p, g, m, v, 1e-2, 0.9, 0.999, 1e-8, step)
    assert torch.allclose(p, pr.detach(), atol=1e-5)
    for T, N in [(16384, 4096), (1000, 264), (77, 8), (4096, 16384)]:
        x = _rand((T, N), cuda_device, seed=23)
        ref = x.double().sum(0)
        out = nv.colsum(x)
        out2 = nv.colsum(x)
        assert torch.equal(out, out2), "column sum must be deterministic"
        assert (out.double() - ref).abs().max().item() <= 1e-3 * max(1.0, ref.abs().max().item()), (T, N)
        acc = out.clone()
        nv.colsum(x, out=acc, accumulate=True)
        assert torch.allclose(acc, 2 * out, rtol=1e-6, atol=1e-4)
    xs = _rand((512, 1024), cuda_device, seed=24)[:, 256:512]             # strided view
    assert (nv.colsum(xs).double() - xs.double().sum(0)).abs().max().item() < 1e-2
    T, N = 300, 512
    res, bias = _rand((T, N), cuda_device, seed=25), torch.randn(N, device=cuda_device)
    out = torch.empty(T, N, dtype=torch.bfloat16, device=cuda_device)
    nv.bias_rows_(out, bias, res)
    assert torch.equal(out, (res.float() + bias).to(torch.bfloat16))
    nv.bias_rows_(out, bias, None)
    assert torch.equal(out, bias.to(torch.bfloat16).expand(T, N))
    nv.bias_rows_(out, None, res)
    assert torch.equal(out, res)


def test_stochastic_rounding_keeps_sub_ulp_updates(cuda_device):
    """bf16 weights have no fp32 master copy: with round-to-nearest an update of 1/16 ulp vanishes entirely;
    with b200_set_update_rounding(seed) the new weight is rounded up with probability 1/16, so the MEAN update is exact.
    Checked on the elementwise SGD kernel and on the SGD epilogue of the dW GEMM; replicas (same seed) agree bit for bit."""
    n = 1 << 20
    ulp = 2.0 ** -7                                         # bf16 ulp at 1.0
    try:
        for kind in ("kernel", "gemm"):
            outs = []
            for seed in (0, 12345, 12345, 999):
                nv.set_update_rounding(seed)
                if kind == "kernel":
                    p = torch.ones(n, dtype=torch.bfloat16, device=cuda_device)
                    g = torch.full((n,), ulp / 16, dtype=torch.float32, device=cuda_device)
                    nv.sgd_(p, g, 1.0)
                else:                                       # W[1024,1024] -= 1 * dy^T x with dy^T x == ulp/16 everywhere (K = 64)
                    p = torch.ones(1024, 1024, dtype=torch.bfloat16, device=cuda_device)
                    dy = torch.full((64, 1024), 2.0 ** -6, dtype=torch.bfloat16, device=cuda_device)
                    x = torch.full((64, 1024), ulp / 16 / 64 * 2.0 ** 6, dtype=torch.bfloat16, device=cuda_device)
                    nv.gemm(dy, 1, x, 1, p, 1024, 1024, 64, nv.EPI_SGD, alpha=1.0)
                torch.cuda.synchronize()
                outs.append(p.float().flatten().clone())
            rn, sr_a, sr_b, sr_c = outs
            assert bool((rn == 1.0).all()), "round-to-nearest drops a 1/16-ulp update (the limitation being addressed)"
            frac = (sr_a < 1.0).float().mean().item()          # 1 - ulp/16 lies between 1 - ulp/2 (next bf16 below 1) and 1
            assert abs(frac - 0.125) < 0.01, (kind, frac)       # below 1.0 the spacing is ulp/2: P(round down) = (ulp/16)/(ulp/2)
            assert abs((1.0 - sr_a.mean().item()) - ulp / 16) < 0.1 * ulp / 16, "mean update must equal the exact update"
            assert torch.equal(sr_a, sr_b), "same seed -> identical rounding (replicas stay in sync)"
            assert not torch.equal(sr_a, sr_c), "a different seed rounds different elements"
    finally:
        nv.set_update_rounding(0)


@pytest.mark.parametrize("layout", ["tn", "nn", "nt"])
@pytest.mark.parametrize("M,N,K", [(256, 512, 256), (200, 264, 192), (1024, 768, 2048)])
def test_gemm_f32_split_operands(cuda_device, layout, M, N, K):
    """fp32 GEMM on the bf16 tensor cores (3 planes per operand, 6 plane products, fp32 accumulate): error against an fp64
    oracle must be at fp32 level - orders of magnitude below a bf16 or TF32 GEMM - in every operand layout."""
    g = torch.Generator(device="cpu").manual_seed(3)
    a = torch.randn(M, K, generator=g).to(cuda_device)
    b = (torch.randn(N, K, generator=g) * 0.1).to(cuda_device)
    ref = a.double() @ b.double().T
    out = torch.empty(M, N, device=cuda_device)
    if layout == "tn":
        nv.gemm_f32(nv.split3(a), 0, nv.split3(b), 0, out, M, N, K)
    elif layout == "nn":
        nv.gemm_f32(nv.split3(a), 0, nv.split3(b.T.contiguous()), 1, out, M, N, K)
    else:
        nv.gemm_f32(nv.split3(a.T.contiguous()), 1, nv.split3(b.T.contiguous()), 1, out, M, N, K)
    torch.cuda.synchronize()
    scale = (a.double().abs() @ b.double().abs().T).max().item()          # sum_k |a||b|: the natural error scale of a dot product
    err = (out.double() - ref).abs().max().item()
    assert err <= 4e-6 * scale, (layout, err, scale)                       # bf16 would be ~4e-3, TF32 ~5e-4
    planes = nv.split3(a)
    assert (planes.double().sum(0) - a.double()).abs().max().item() <= 2 ** -23 * a.abs().max().item()


def test_gemm_f32_epilogues(cuda_device):
    M, N, K = 384, 512, 320
    g = torch.Generator(device="cpu").manual_seed(4)
    x, w = torch.randn(M, K, generator=g).to(cuda_device), (torch.randn(N, K, generator=g) * 0.1).to(cuda_device)
    aux, bias = torch.randn(M, N, generator=g).to(cuda_device), torch.randn(N, generator=g).to(cuda_device)
    ref = (x.double() @ w.double().T)
    X, W = nv.split3(x), nv.split3(w)
    out, out2 = torch.empty(M, N, device=cuda_device), torch.empty(M, N, device=cuda_device)
    tol = 1e-4

    def close(o, r, what):
        assert (o.double() - r).abs().max().item() <= tol * max(1.0, r.abs().max().item()), what

    nv.gemm_f32(X, 0, W, 0, out, M, N, K, nv.EPI_RELU, bias=bias)
    close(out, torch.relu(ref + bias.double()), "relu+bias")
    nv.gemm_f32(X, 0, W, 0, out, M, N, K, nv.EPI_GELU, out2=out2)
    close(out, torch.nn.functional.gelu(ref), "gelu")
    close(out2, ref, "gelu pre-activation")
    nv.gemm_f32(X, 0, W, 0, out, M, N, K, nv.EPI_RELU_BWD, aux=aux)
    close(out, ref * (aux.double() > 0), "relu_bwd")
    nv.gemm_f32(X, 0, W, 0, out, M, N, K, nv.EPI_GELU_BWD, aux=aux)
    a64 = aux.double().requires_grad_(True)
    torch.nn.functional.gelu(a64).sum().backward()
    close(out, ref * a64.grad, "gelu_bwd")
    nv.gemm_f32(X, 0, W, 0, out, M, N, K, nv.EPI_ADD, aux=aux, bias=bias)
    close(out, ref + aux.double() + bias.double(), "add")
    wgt = torch.randn(M, N, generator=g).to(cuda_device)
    w0 = wgt.double().clone()
    nv.gemm_f32(X, 0, W, 0, wgt, M, N, K, nv.EPI_SGD, alpha=0.05)
    close(wgt, w0 - 0.05 * ref, "sgd")
    nv.gemm_f32(X, 0, W, 0, out, M, N, K, nv.EPI_STORE, alpha=0.5)
    close(out, 0.5 * ref, "store*alpha")


@pytest.mark.parametrize("R,C", [(16384, 4096), (200, 264), (66, 2), (64, 64)])
def test_native_transpose(cuda_device, R, C):
    x = _rand((R, C), cuda_device, seed=51)
    assert torch.equal(nv.transpose(x), x.t().contiguous())
    xs = _rand((R, 2 * C + 4), cuda_device, seed=52)[:, 2:C + 2]                  # strided view with an odd-ish offset
    assert torch.equal(nv.transpose(xs), xs.t().contiguous())
