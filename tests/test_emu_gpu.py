"""Multi-rank numerics of the fused GEMM+collective kernels on ONE B200: every emulated rank runs its real kernel body as
a slice of one cooperative launch (``parallel/emu.py``), against an fp64 oracle of the bf16 inputs, tile by tile.

What this covers that the world-of-one tests cannot: owner maps with several owners (interleaved and row-sharded), the
rotated tile walks, inbox slots indexed by source rank, both inbox parities and flag reuse over consecutive calls, the
broadcast of reduced tiles into every replica, ready-flag delivery, the sharded Adam state, and the all-gather+GEMM
kernel's pull order / per-block readiness in both the row-sharded and the K-sharded form - at 2, 4 and 8 ranks and at
the benchmark's tile counts (2048 tiles).  The reference's counterparts are its NCCL calls (``train_ffns.py:165``,
``:203``, ``:255-256``, ``:303``, ``:309``) and its own DDP==FSDP check (``:386-391``)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributed_llm_code_samples_b200.ops import native as nv  # noqa: E402

TM, TN = 128, 256


def _randn(shape, dev, scale=1.0, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    return (torch.randn(shape, generator=g, device=dev) * scale).to(torch.bfloat16)


def _tile_err(out: torch.Tensor, ref: torch.Tensor) -> float:
    """max over 128x256 tiles of (max abs error in the tile) / (max |ref| in the tile)."""
    M, N = ref.shape
    worst = 0.0
    d = (out.double() - ref).abs()
    for m0 in range(0, M, TM):
        dm, rm = d[m0:m0 + TM], ref[m0:m0 + TM].abs()
        # column blocks in one shot
        nb = (N + TN - 1) // TN
        pad = nb * TN - N
        if pad:
            dm = torch.nn.functional.pad(dm, (0, pad))
            rm = torch.nn.functional.pad(rm, (0, pad))
        e = dm.reshape(dm.shape[0], nb, TN).amax(dim=(0, 2))
        s = rm.reshape(rm.shape[0], nb, TN).amax(dim=(0, 2)).clamp_min(1e-30)
        worst = max(worst, float((e / s).max()))
    return worst


def _oracle_sum(A, a_major, B, b_major):
    """fp64 sum over ranks of A_r . B_r (operand conventions of b200_gemm_bf16)."""
    tot = None
    for a, b in zip(A, B):
        a64 = a.double() if a_major == nv.MAJOR_K else a.double().t()
        b64 = b.double().t() if b_major == nv.MAJOR_K else b.double()
        p = a64 @ b64
        tot = p if tot is None else tot + p
    return tot


def _channel(world, M, N, owner_mode, dev):
    from distributed_llm_code_samples_b200.parallel.emu import EmuFused
    lib = nv.load()
    cap = int(lib.b200_fused_owned_tiles(M, N, world, owner_mode))
    tiles = ((M + TM - 1) // TM) * ((N + TN - 1) // TN)
    return EmuFused(world, cap, tiles, dev)


@pytest.fixture(autouse=True)
def _reset_variant():
    yield
    nv.load().b200_fused_config(4, -1)


@pytest.mark.parametrize("side,overlap", [(4, 0), (2, 0), (4, 1), (2, 1)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_emu_allreduce_sum_interleaved(cuda_device, world, side, overlap):
    """Row-parallel GEMM + all-reduce (MP form, train_ffns.py:57+303): TN operands, interleaved owners, broadcast; three
    calls on one channel (inbox parities 1, 0, 1), ragged M."""
    from distributed_llm_code_samples_b200.parallel.emu import FOP_SUM, OWN_INTERLEAVED
    nv.load().b200_fused_config(side, overlap)
    M, N, K = 2000, 1024, 320
    ch = _channel(world, M, N, OWN_INTERLEAVED, cuda_device)
    for it in range(3):
        A = [_randn((M, K), cuda_device, seed=10 * it + r) for r in range(world)]
        B = [_randn((N, K), cuda_device, 0.1, seed=100 + 10 * it + r) for r in range(world)]
        out = [torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
        ch.gemm(A, nv.MAJOR_K, B, nv.MAJOR_K, M, N, K, out, OWN_INTERLEAVED, FOP_SUM, True)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        ref = _oracle_sum(A, nv.MAJOR_K, B, nv.MAJOR_K)
        for r in range(world):
            assert torch.isfinite(out[r].float()).all(), (it, r)
            assert _tile_err(out[r], ref) <= 1e-2, (it, r, _tile_err(out[r], ref))
        for r in range(1, world):
            assert torch.equal(out[r], out[0]), "replicas must be bit-identical (fixed-order fp32 sum)"


@pytest.mark.parametrize("side,overlap", [(4, 0), (2, 0), (4, 1)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_emu_reduce_scatter_sum_row_sharded(cuda_device, world, side, overlap):
    """Row-parallel GEMM + reduce-scatter over token shards (sequence-parallel MP): NN operands (dx = dh W1), row-sharded
    owners with rotated tile walks, no broadcast."""
    from distributed_llm_code_samples_b200.parallel.emu import FOP_SUM, OWN_ROW_SHARDED
    nv.load().b200_fused_config(side, overlap)
    M, N, K = 128 * world * 3, 768, 448
    ch = _channel(world, M, N, OWN_ROW_SHARDED, cuda_device)
    for it in range(3):
        A = [_randn((M, K), cuda_device, seed=20 * it + r) for r in range(world)]
        B = [_randn((K, N), cuda_device, 0.1, seed=300 + 20 * it + r) for r in range(world)]
        out = [torch.full((M // world, N), float("nan"), dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
        ch.gemm(A, nv.MAJOR_K, B, nv.MAJOR_MN, M, N, K, out, OWN_ROW_SHARDED, FOP_SUM, False)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        ref = _oracle_sum(A, nv.MAJOR_K, B, nv.MAJOR_MN)
        for r in range(world):
            sl = ref[r * (M // world):(r + 1) * (M // world)]
            assert _tile_err(out[r], sl) <= 1e-2, (it, r, _tile_err(out[r], sl))


@pytest.mark.parametrize("side", [4, 2])
@pytest.mark.parametrize("world,mode", [(2, "ddp"), (4, "ddp"), (8, "ddp"), (2, "fsdp"), (4, "fsdp"), (8, "fsdp")])
def test_emu_dw_sgd(cuda_device, world, mode, side):
    """dW GEMM + all-reduce + SGD with the new weight stored into every replica (DDP, train_ffns.py:66/68+165+172), and
    dW GEMM + reduce-scatter + sharded SGD (FSDP, :255-259): NT operands (K = tokens), two consecutive steps."""
    from distributed_llm_code_samples_b200.parallel.emu import FOP_SGD, OWN_INTERLEAVED, OWN_ROW_SHARDED
    nv.load().b200_fused_config(side, -1)
    T, Dout, Din = 640, 128 * world * 2, 512
    lr = 0.01
    om = OWN_INTERLEAVED if mode == "ddp" else OWN_ROW_SHARDED
    ch = _channel(world, Dout, Din, om, cuda_device)
    W0 = _randn((Dout, Din), cuda_device, 0.5, seed=7)
    if mode == "ddp":
        W = [W0.clone() for _ in range(world)]
    else:
        W = [W0[r * (Dout // world):(r + 1) * (Dout // world)].clone() for r in range(world)]
    ref = W0.double()
    for it in range(2):
        dy = [_randn((T, Dout), cuda_device, seed=40 + 10 * it + r) for r in range(world)]
        x = [_randn((T, Din), cuda_device, seed=400 + 10 * it + r) for r in range(world)]
        ch.gemm(dy, nv.MAJOR_MN, x, nv.MAJOR_MN, Dout, Din, T, W, om, FOP_SGD, mode == "ddp", alpha=lr)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        grad = _oracle_sum(dy, nv.MAJOR_MN, x, nv.MAJOR_MN)
        # the oracle follows the kernel's bf16 weight grid: next step starts from the stored (rounded) weights
        ref = ref - lr * grad
        full = W[0] if mode == "ddp" else torch.cat(W, 0)
        upd_err = ((full.double() - ref).abs().max() / (lr * grad.abs().max())).item()
        assert upd_err <= 2.5e-2 + 0.5 * 2 ** -8 * float(ref.abs().max()) / (lr * float(grad.abs().max())), (it, upd_err)
        if mode == "ddp":
            for r in range(1, world):
                assert torch.equal(W[r], W[0])
        ref = full.double()


@pytest.mark.parametrize("world,mode", [(2, "ddp"), (8, "ddp"), (4, "fsdp")])
def test_emu_dw_adam(cuda_device, world, mode):
    """Adam inside the owner's reduction, moments sharded by tile owner, two steps (bias correction + stored moments)."""
    from distributed_llm_code_samples_b200.parallel.emu import FOP_SGD, OWN_INTERLEAVED, OWN_ROW_SHARDED
    lib = nv.load()
    T, Dout, Din = 512, 128 * world, 768
    lr, b1, b2, eps = 1e-3, 0.9, 0.999, 1e-8
    om = OWN_INTERLEAVED if mode == "ddp" else OWN_ROW_SHARDED
    ch = _channel(world, Dout, Din, om, cuda_device)
    n_state = int(lib.b200_fused_adam_state_bytes(ch.owned_cap)) // 4
    state = [(torch.zeros(n_state, device=cuda_device), torch.zeros(n_state, device=cuda_device)) for _ in range(world)]
    W0 = _randn((Dout, Din), cuda_device, 0.5, seed=9)
    W = [W0.clone() for _ in range(world)] if mode == "ddp" else [W0[r * (Dout // world):(r + 1) * (Dout // world)].clone() for r in range(world)]
    ref, m, v = W0.double(), torch.zeros_like(W0, dtype=torch.float64), torch.zeros_like(W0, dtype=torch.float64)
    for step in (1, 2):
        dy = [_randn((T, Dout), cuda_device, seed=60 + 10 * step + r) for r in range(world)]
        x = [_randn((T, Din), cuda_device, seed=600 + 10 * step + r) for r in range(world)]
        ch.gemm(dy, nv.MAJOR_MN, x, nv.MAJOR_MN, Dout, Din, T, W, om, FOP_SGD, mode == "ddp", adam_state=state, lr=lr, step=step,
                beta1=b1, beta2=b2, eps=eps)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        g = _oracle_sum(dy, nv.MAJOR_MN, x, nv.MAJOR_MN)
        m = b1 * m + (1 - b1) * g
        v = b2 * v + (1 - b2) * g * g
        ref = ref - lr * (m / (1 - b1 ** step)) / ((v / (1 - b2 ** step)).sqrt() + eps)
        full = W[0] if mode == "ddp" else torch.cat(W, 0)
        # every weight moves by ~lr; allow the bf16 grid of the weight (|W| <= ~2.5 -> ulp 2^-7) plus 5 % of the step
        assert (full.double() - ref).abs().max().item() <= 2 ** -7 + 0.05 * lr, step
        ref = full.double()


def test_emu_benchmark_shapes_8_ranks(cuda_device):
    """The benchmark's dW shapes at 8 ranks (2048 tiles of 128x256, K = 16384 tokens; ~14 owned tiles per CTA) as plain
    sums (op = SUM exposes the summed GRADIENT, the sharpest check), interleaved + broadcast (DDP) and row-sharded (FSDP);
    then the MP all-reduce shape (16384 x 8192, K = 3584: 4096 tiles)."""
    from distributed_llm_code_samples_b200.parallel.emu import FOP_SUM, OWN_INTERLEAVED, OWN_ROW_SHARDED
    world = 8
    for (M, N, K, a_major, b_major, om, bc) in [
        (16384, 4096, 16384, nv.MAJOR_MN, nv.MAJOR_MN, OWN_INTERLEAVED, True),
        (4096, 16384, 16384, nv.MAJOR_MN, nv.MAJOR_MN, OWN_ROW_SHARDED, False),
        (16384, 8192, 3584, nv.MAJOR_K, nv.MAJOR_K, OWN_INTERLEAVED, True),
    ]:
        ch = _channel(world, M, N, om, cuda_device)
        if a_major == nv.MAJOR_MN:
            A = [_randn((K, M), cuda_device, 0.05, seed=r) for r in range(world)]
            B = [_randn((K, N), cuda_device, 0.05, seed=50 + r) for r in range(world)]
        else:
            A = [_randn((M, K), cuda_device, 0.05, seed=r) for r in range(world)]
            B = [_randn((N, K), cuda_device, 0.05, seed=50 + r) for r in range(world)]
        rows = M if bc else M // world
        out = [torch.full((rows, N), float("nan"), dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
        for it in range(2):                                        # both inbox parities at full size
            ch.gemm(A, a_major, B, b_major, M, N, K, out, om, FOP_SUM, bc)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        ref = _oracle_sum(A, a_major, B, b_major)
        if bc:
            assert _tile_err(out[0], ref) <= 1e-2, (M, N, K, _tile_err(out[0], ref))
            for r in range(1, world):
                assert torch.equal(out[r], out[0])
        else:
            for r in range(world):
                sl = ref[r * rows:(r + 1) * rows]
                assert _tile_err(out[r], sl) <= 1e-2, (M, N, K, r)
        del ch, A, B, out, ref
        torch.cuda.empty_cache()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("epi", ["store", "relu_mask", "gelu"])
def test_emu_gather_gemm_row_sharded(cuda_device, world, epi):
    """all-gather -> column-parallel GEMM in one kernel (MP spf forward; FSDP forward with the weight as A): A is sharded
    by rows; every rank ends with the complete A and with epilogue(A_full . B_r^T).  Three calls (flag reuse)."""
    from distributed_llm_code_samples_b200.parallel.emu import EmuGather
    M, N, K = 128 * world * 2, 640, 448          # N % 256 == 128: the mask epilogue's half-empty last column tile
    kind = {"store": nv.EPI_STORE, "relu_mask": nv.EPI_RELU_MASK, "gelu": nv.EPI_GELU}[epi]
    ch = EmuGather(world, M, cuda_device)
    Ms = M // world
    for it in range(3):
        full = _randn((M, K), cuda_device, seed=70 + it)
        bufs = []
        for r in range(world):
            b = torch.full((M, K), float("nan"), dtype=torch.bfloat16, device=cuda_device)
            b[r * Ms:(r + 1) * Ms] = full[r * Ms:(r + 1) * Ms]
            bufs.append(b)
        B = [_randn((N, K), cuda_device, 0.1, seed=700 + 10 * it + r) for r in range(world)]
        out = [torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
        out2 = None
        if epi == "relu_mask":
            out2 = [nv.relu_mask_empty(M, N, cuda_device).zero_() for _ in range(world)]
        elif epi == "gelu":
            out2 = [torch.empty((M, N), dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
        ch.gemm(bufs, nv.MAJOR_K, B, nv.MAJOR_K, out, M, N, K, kind, out2=out2)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        for r in range(world):
            assert torch.equal(bufs[r], full), (it, r, "gathered A incomplete")
            h = full.double() @ B[r].double().t()
            ref = h if epi == "store" else (torch.relu(h) if epi == "relu_mask" else torch.nn.functional.gelu(h))
            assert _tile_err(out[r], ref) <= 1e-2, (it, r)
            if epi == "relu_mask":
                got = nv.unpack_relu_mask(out2[r], M, N)
                assert torch.equal(got, out[r] > 0), (it, r, "sign mask disagrees with the activation")


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emu_gather_gemm_k_sharded(cuda_device, world):
    """The K-sharded form (FSDP backward: the weight's sharded dimension is the contracted one): A stored [K, M], sharded
    along K; every tile's k-loop starts at the rank's own k-range and waits per k-block."""
    from distributed_llm_code_samples_b200.parallel.emu import EmuGather
    K, M, N = 128 * world * 2, 512, 384
    ch = EmuGather(world, K, cuda_device)
    Ks = K // world
    for it in range(2):
        full = _randn((K, M), cuda_device, seed=80 + it)
        bufs = []
        for r in range(world):
            b = torch.full((K, M), float("nan"), dtype=torch.bfloat16, device=cuda_device)
            b[r * Ks:(r + 1) * Ks] = full[r * Ks:(r + 1) * Ks]
            bufs.append(b)
        B = [_randn((N, K), cuda_device, 0.1, seed=800 + 10 * it + r) for r in range(world)]
        out = [torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
        ch.gemm(bufs, nv.MAJOR_MN, B, nv.MAJOR_K, out, M, N, K, nv.EPI_STORE)
        torch.cuda.synchronize()
        assert nv.device_error()[0] == 0, nv.device_error()
        for r in range(world):
            assert torch.equal(bufs[r], full), (it, r)
            ref = full.double().t() @ B[r].double().t()
            assert _tile_err(out[r], ref) <= 1e-2, (it, r)
