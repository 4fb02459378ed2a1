"""Fused GEMM+collective kernel and native collectives in the degenerate world of ONE rank (single B200): the
whole push -> flag -> reduce -> op -> store path runs, with the rank pushing to its own inbox."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from distributed_llm_code_samples_b200.ops import native as nv  # noqa: E402


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16).to(dev)


@pytest.fixture(scope="module")
def sw(cuda_device):
    from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
    w = SymmWorld(0, 1, cuda_device)
    yield w
    w.close()


@pytest.mark.parametrize("owner_mode", [0, 1])
@pytest.mark.parametrize("M,N,K", [(256, 512, 192), (1024, 768, 2048), (384, 264, 520)])
def test_fused_sum_world1(cuda_device, sw, owner_mode, M, N, K):
    from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles
    if owner_mode == 1 and M % 128:
        pytest.skip("row-sharded owners need M % (128*world) == 0")
    a, w = _rand((M, K), cuda_device, seed=1), _rand((N, K), cuda_device, 0.1, seed=2)
    chan = FusedChannel(sw, _owned_cap([(M, N)], 1, owner_mode), _tiles(M, N))
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda_device)
    tab = (C.c_void_p * 1)(out.data_ptr())
    ref = a.float() @ w.float().T
    for it in range(3):                               # parity 1, 0, 1: both inbox halves and flag reuse
        out.fill_(float("nan"))
        chan.gemm(a, 0, w, 0, M, N, K, tab, N, 0, owner_mode, 0, True)
        torch.cuda.synchronize()
        err = (out.float() - ref).abs().max().item()
        assert err <= 2e-2 * max(1.0, ref.abs().max().item()), (it, err)
    assert nv.device_error()[0] == 0


def test_fused_sgd_nt_world1(cuda_device, sw):
    """dW = dy^T x with the optimizer step applied by the owner: W -= lr * dW (MN-major operands, K = tokens)."""
    from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles
    T, Dout, Din = 640, 512, 256
    dy, x = _rand((T, Dout), cuda_device, seed=3), _rand((T, Din), cuda_device, seed=4)
    W = _rand((Dout, Din), cuda_device, 0.5, seed=5)
    W0 = W.float().clone()
    chan = FusedChannel(sw, _owned_cap([(Dout, Din)], 1, 1), _tiles(Dout, Din))
    tab = (C.c_void_p * 1)(W.data_ptr())
    chan.gemm(dy, 1, x, 1, Dout, Din, T, tab, Din, 0, 1, 1, False, alpha=0.01)
    torch.cuda.synchronize()
    ref = W0 - 0.01 * (dy.float().T @ x.float())
    assert (W.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())


def test_native_collectives_world1(cuda_device, sw):
    from distributed_llm_code_samples_b200.parallel.symm import SymmComm
    comm = SymmComm(sw, staging_bytes=8 << 20)
    t = _rand((4096,), cuda_device, seed=6)
    t0 = t.clone()
    comm.all_reduce(t)
    comm.barrier()
    big = _rand((1 << 20,), cuda_device, seed=7)
    b0 = big.clone()
    h = comm.all_reduce(big, async_op=True)
    h.wait()
    out = torch.empty_like(big)
    comm.all_gather_into(out, big)
    rs = torch.empty_like(big)
    comm.reduce_scatter_into(rs, big)
    torch.cuda.synchronize()
    assert torch.equal(t, t0) and torch.equal(big, b0) and torch.equal(out, b0) and torch.equal(rs, b0)
    assert nv.device_error()[0] == 0
