"""Dry run of the fused (GPU-only) trainers on CPU.

The peer-memory trainers in parallel/fused.py never execute in the CPU suite, so a Python slip or a mis-ordered C-ABI
argument would first show up on the GPU box.  Here they run against
  * a fake symmetric world (CPU memory, every "peer" pointer = the local block),
  * the REAL native library: each call goes through ctypes marshalling and the entry point's host-side validation
    (shapes, strides, alignment, inbox / flag capacities, owner maps) and only then fails at its first CUDA call,
    which this harness tolerates.  Any B200_ERR_ARG (-1) is a test failure.
The numbers the steps produce are meaningless (no kernel runs); what is checked is that every call is well-formed."""
import ctypes as C

import pytest
import torch

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import get_ops, native as nv
from distributed_llm_code_samples_b200.parallel import fused, symm
from distributed_llm_code_samples_b200.parallel.base import fsdp_shard, tp_shard

# On a GPU box the "first CUDA call" would really launch kernels on CPU pointers: this harness is for GPU-less boxes only.
pytestmark = pytest.mark.skipif(torch.cuda.is_available(), reason="dry run is meant for boxes without a GPU")

D, FF, L, T, WORLD = 256, 1024, 2, 512, 2


class _FakeBuffer(symm.SymmBuffer):
    def __init__(self, world, nbytes):          # noqa: super().__init__ would cudaMalloc
        self.world, self.nbytes = world, nbytes
        self._store = torch.zeros(nbytes + 512, dtype=torch.uint8)
        pad = (-self._store.data_ptr()) % 256
        self._base = self._store[pad: pad + nbytes]
        self.ptr = self._base.data_ptr()
        self.peer_ptrs = [self.ptr] * world.world
        self.peers_c = (C.c_void_p * world.world)(*self.peer_ptrs)
        self._opened, self._cursor = [], 0
        self.mc_ptr = None


class _FakeWorld:
    def __init__(self, rank, world):
        self.rank, self.world, self.device, self.group, self.buffers = rank, world, torch.device("cpu"), None, []

    def multicast_supported(self):
        return False

    def alloc(self, nbytes, multicast=False):
        b = _FakeBuffer(self, (nbytes + 255) // 256 * 256)
        self.buffers.append(b)
        return b

    def host_barrier(self):
        pass


class _Comm:
    def __init__(self, rank, world):
        self.rank, self.world = rank, world


class _Stream:
    cuda_stream = 0

    def __init__(self, *a, **k):
        pass

    def wait_stream(self, *_):
        pass

    def wait_event(self, *_):
        pass


class _Event:
    def record(self, *_):
        pass


@pytest.fixture
def dry(monkeypatch):
    calls = []

    def check(rc, what):
        calls.append((what, rc))
        if rc == -1:
            raise AssertionError(f"{what}: argument rejected: {nv.load().b200_last_error().decode()}")

    monkeypatch.setattr(nv, "_check", check)
    monkeypatch.setattr(nv, "_stream", lambda: 0)
    monkeypatch.setattr(torch.cuda, "Stream", _Stream)
    monkeypatch.setattr(torch.cuda, "Event", _Event)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: _Stream())
    monkeypatch.setattr(fused, "get_symm_world", lambda comm, device: _FakeWorld(comm.rank, comm.world))
    return calls


def _cfg(**kw):
    base = dict(model_size=D, ffn_dim=FF, layers=L, dtype="bf16", lr=1e-3, world_size=WORLD)
    base.update(kw)
    return TrainConfig(**base)


def _batch():
    g = torch.Generator().manual_seed(0)
    return torch.randn(T, D, generator=g).to(torch.bfloat16), (0.1 * torch.randn(T, D, generator=g)).to(torch.bfloat16)


def _params():
    return [[p.to(torch.bfloat16) for p in l] for l in ffn.init_stack(L, D, FF, torch.Generator().manual_seed(1))]


def _names(calls):
    return [w for w, _ in calls]


@pytest.mark.parametrize("rank", [0, 1])
@pytest.mark.parametrize("optimizer", ["sgd", "adam"])
def test_fused_ddp_calls_are_well_formed(dry, rank, optimizer):
    tr = fused.FusedDDPTrainer(get_ops("torch", "cpu"), _Comm(rank, WORLD), _cfg(optimizer=optimizer), _params(), torch.device("cpu"))
    tr.step(*_batch())
    tr.step(*_batch())
    want = "b200_gemm_allreduce_adam_bf16" if optimizer == "adam" else "b200_gemm_allreduce_bf16"
    assert _names(dry).count(want) == 2 * 2 * L and all(rc != 0 for _, rc in dry), "no GPU: every call stops at its first CUDA call"
    assert tr.chan.seq == 2 * 2 * L and tr.t == 2


@pytest.mark.parametrize("rank", [0, 1])
@pytest.mark.parametrize("optimizer", ["sgd", "adam"])
def test_fused_fsdp_calls_are_well_formed(dry, rank, optimizer):
    shards = [[fsdp_shard(p, rank, WORLD) for p in l] for l in _params()]
    tr = fused.FusedFSDPTrainer(get_ops("torch", "cpu"), _Comm(rank, WORLD), _cfg(optimizer=optimizer), shards, torch.device("cpu"))
    tr.step(*_batch())
    names = _names(dry)
    want = "b200_gemm_allreduce_adam_bf16" if optimizer == "adam" else "b200_gemm_allreduce_bf16"
    assert names.count(want) == 2 * L
    assert names.count("b200_all_gather_into_dma") == 2 * (2 * L - 1), "prefetch: L forward + L-1 backward layers, two weights each (:244-249)"
    tr.full_params()


@pytest.mark.parametrize("rank", [0, 1])
@pytest.mark.parametrize("mp_mode", ["ar", "sp", "spf"])
@pytest.mark.parametrize("residual,act,optimizer", [(False, "relu", "sgd"), (True, "gelu", "adam")])
def test_fused_mp_calls_are_well_formed(dry, rank, mp_mode, residual, act, optimizer):
    shards = [[tp_shard(p, j, rank, WORLD).contiguous() for j, p in enumerate(l)] for l in _params()]
    cfg = _cfg(mp_mode=mp_mode, residual=residual, activation=act, optimizer=optimizer)
    tr = fused.FusedMPTrainer(get_ops("torch", "cpu"), _Comm(rank, WORLD), cfg, shards, torch.device("cpu"))
    y = tr.step(*_batch())
    assert y.shape == (T, D)
    names = _names(dry)
    assert names.count("b200_gemm_allreduce_bf16") == 2 * L - 1, "one fused row-parallel GEMM per layer forward, L-1 backward (:303, :309)"
    if mp_mode == "spf":
        assert names.count("b200_gather_gemm_bf16") == 2 * L and names.count("b200_all_gather") == 1
    if mp_mode == "sp":
        assert names.count("b200_all_gather_into") == L + 2 * L + 1


def test_native_collective_wrappers_are_well_formed(dry):
    """SymmComm (--comm native): in-place (symmetric operand) and staged (ordinary tensor) forms of every collective."""
    sw = _FakeWorld(1, WORLD)
    comm = symm.SymmComm(sw, staging_bytes=1 << 20)
    inplace = sw.alloc(1 << 16).tensor(0, (4096,), torch.bfloat16)
    comm.barrier()
    for t in (inplace, torch.zeros(1000, dtype=torch.bfloat16), torch.zeros(1000, dtype=torch.float32)):
        comm.all_reduce(t)
    comm.all_gather_into(torch.zeros(2 * 4096, dtype=torch.bfloat16), inplace)
    comm.all_gather_into(torch.zeros(2 * 512, dtype=torch.bfloat16), torch.zeros(512, dtype=torch.bfloat16))
    comm.reduce_scatter_into(torch.zeros(2048, dtype=torch.bfloat16), inplace)
    comm.reduce_scatter_into(torch.zeros(256, dtype=torch.float32), torch.zeros(512, dtype=torch.float32))
    h = comm.all_reduce(inplace, async_op=True)
    h.wait()
    names = _names(dry)
    assert names.count("b200_all_reduce") == 4 and names.count("b200_all_gather_into") == 2 and names.count("b200_reduce_scatter") == 2
    assert names.count("b200_barrier") == 1 and all(rc != 0 for _, rc in dry)
    assert comm.chans[0].epoch + comm.chans[1].epoch == 9, "every collective consumed exactly one epoch of its lane"
    assert comm.chans[1].epoch >= 1, "the async call ran on the side-stream lane (its own flag block and staging half)"


def test_sharded_input_staging_is_well_formed(dry, monkeypatch):
    """utils/data.py::ShardedPrefetcher: double-buffered staging of 1/world of the rows + two native all-gathers per batch."""
    from distributed_llm_code_samples_b200.utils.data import ShardedPrefetcher
    monkeypatch.setattr(torch.Tensor, "is_pinned", lambda self: True)
    batches = [(torch.full((T, D), float(i)).to(torch.bfloat16), torch.full((T, D), -float(i)).to(torch.bfloat16)) for i in range(4)]
    sw = _FakeWorld(1, WORLD)
    pf = ShardedPrefetcher(iter(batches), torch.device("cpu"), sw, 1, WORLD)
    seen = 0
    for i, (x, dy) in enumerate(pf):
        assert x.shape == (T, D) and dy.shape == (T, D) and pf.last_bytes == 2 * (T // WORLD) * D * 2
        # the staging copy itself is real even here: my rows of batch i+1 are already in the other slot
        seen += 1
    assert seen == 4 and _names(dry).count("b200_all_gather_into") == 8
    ts = T // WORLD
    last = pf.stage.tensor(pf.slot_bytes[0] * 0 + (3 % 2) * sum(pf.slot_bytes), (ts, D), torch.bfloat16)
    assert torch.equal(last, batches[3][0][ts:2 * ts]), "slot 1 holds rank 1's rows of the last batch"


@pytest.mark.parametrize("rank", [0, 1])
@pytest.mark.parametrize("act,residual,optimizer", [("relu", False, "sgd"), ("gelu", True, "adam")])
def test_fused_fsdp_with_in_kernel_forward_gather_is_well_formed(dry, rank, act, residual, optimizer):
    from distributed_llm_code_samples_b200.parallel.fused_fsdp_t import FusedFSDPGatherTrainer
    shards = [[fsdp_shard(p, rank, WORLD) for p in l] for l in _params()]
    cfg = _cfg(activation=act, residual=residual, optimizer=optimizer, layers=4)
    four = shards + [[p.clone() for p in l] for l in shards]
    tr = FusedFSDPGatherTrainer(get_ops("torch", "cpu"), _Comm(rank, WORLD), cfg, four, torch.device("cpu"))
    y = tr.step(*_batch())
    assert y.shape == (T, D)
    names = _names(dry)
    # forward: both weights of all 4 layers; backward: layers 3 and 2 are still complete from the forward pass, layer 1
    # gathers W2 and W1 (K-sharded form), layer 0 gathers W2 only (its input gradient is not needed, :88)
    assert names.count("b200_gather_gemm_bf16") == 2 * 4 + 2 + 1
    assert names.count("b200_gemm_bf16") == 2 + 2, "dh^T and dx^T of the two layers whose weights the forward pass left behind"
    assert names.count("b200_gemm_allreduce_adam_bf16" if optimizer == "adam" else "b200_gemm_allreduce_bf16") == 2 * 4
    assert names.count("b200_all_gather_into_dma") == 0, "no copy-engine prefetch is left"
