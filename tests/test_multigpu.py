"""Multi-GPU tests (>= 2 B200 on one node; run with ``gpurun --gpus 2``): symmetric memory, native collectives
vs closed-form sums (the reference's test_nccl.py checks), fused trainers vs the NCCL trainers."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")]

import os

WORLD = min(torch.cuda.device_count(), 8) if torch.cuda.is_available() else 0
# Tests of code that has not run on hardware yet are opt-in until validated (none at present: round 2's first 2-GPU call
# ran everything round 1 had queued - log in profiles/tests/r2_call1_2gpu/).
queued = pytest.mark.skipif(os.environ.get("B200_QUEUED_TESTS") != "1", reason="queued for the next GPU round (set B200_QUEUED_TESTS=1)")
DM = max(256, 128 * max(WORLD, 1))


CHECKS = ["p2p_write", "ar_128", "ag_128", "rs_128", "ar_64k", "ag_64k", "rs_64k", "ar_4m", "ag_4m", "rs_4m", "async_side_stream", "no_device_error"]


def _coll_worker(rank, world, dev, results):
    from distributed_llm_code_samples_b200.ops import native as nv
    from distributed_llm_code_samples_b200.parallel.symm import SymmComm, SymmWorld
    sw = SymmWorld(rank, world, dev)
    comm = SymmComm(sw, staging_bytes=64 << 20)
    res = {}
    # P2P visibility (test_mp_barrier_gpus.py): every rank mutates ITS symmetric block; after the device barrier
    # every other rank reads those bytes straight out of the peer mapping
    buf = sw.alloc(4096)
    mine = buf.tensor(0, (512,), torch.float32)
    mine.fill_(float(rank + 1))
    comm.barrier()
    seen = torch.empty(512 * world, dtype=torch.float32, device=dev)
    comm.all_gather_into(seen, mine)
    torch.cuda.synchronize()
    expect = torch.arange(1, world + 1, dtype=torch.float32).repeat_interleave(512)
    res["p2p_write"] = torch.equal(seen.cpu(), expect)
    # collectives vs closed form (test_nccl.py:10,23,36), small (one-shot) and large (two-shot)
    for tag, n in (("128", 128), ("64k", 1 << 16), ("4m", (1 << 22) + 8 * world)):
        n = max(n // (8 * world) * (8 * world), 8 * world)
        g = torch.Generator().manual_seed(100 + n)
        allx = [(torch.randn(n, generator=g)).to(torch.bfloat16) for _ in range(world)]
        total = sum(t.float() for t in allx)
        x = allx[rank].to(dev)
        comm.all_reduce(x)
        torch.cuda.synchronize()
        res["ar_" + tag] = bool((x.float().cpu() - total).abs().max() <= 0.05 * world)
        shard = allx[rank][: n // world].contiguous().to(dev)
        out = torch.empty(n, dtype=torch.bfloat16, device=dev)
        comm.all_gather_into(out, shard)
        torch.cuda.synchronize()
        res["ag_" + tag] = torch.equal(out.cpu(), torch.cat([t[: n // world] for t in allx]))
        rs = torch.empty(n // world, dtype=torch.bfloat16, device=dev)
        comm.reduce_scatter_into(rs, allx[rank].to(dev))
        torch.cuda.synchronize()
        sl = total[rank * (n // world): (rank + 1) * (n // world)]
        res["rs_" + tag] = bool((rs.float().cpu() - sl).abs().max() <= 0.05 * world)
    # async handle + side stream (test_torch_distributed.py:13-20): all-reduce of ones, then +100 -> world + 100
    t = torch.ones(8, dtype=torch.bfloat16, device=dev)
    h = comm.all_reduce(t, async_op=True)
    h.wait()
    s = torch.cuda.Stream(device=dev)
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        t.add_(100)
    torch.cuda.synchronize()
    res["async_side_stream"] = bool((t.float() == world + 100).all())
    res["no_device_error"] = nv.device_error()[0] == 0
    for i, k in enumerate(CHECKS):
        results[rank, i] = 1.0 if res.get(k, False) else 0.0
    sw.close()


def test_native_collectives_and_p2p():
    from distributed_llm_code_samples_b200.parallel.launcher import launch
    results = torch.zeros(WORLD, len(CHECKS)).share_memory_()
    launch(WORLD, _coll_worker, (results,), backend="nccl", device="cuda")
    failed = [(r, CHECKS[i]) for r in range(WORLD) for i in range(len(CHECKS)) if results[r, i] != 1.0]
    assert not failed, failed


def _grad_worker(rank, world, dev, results):
    """Every rank computes its partial product; the fused kernel's SUM must equal the fp64 sum of all ranks' products, per
    128x256 tile, for interleaved + broadcast (DDP form) and row-sharded + local (FSDP form) owners, 3 calls each (both inbox
    parities), NT operands like the dW GEMMs."""
    import ctypes as C
    from distributed_llm_code_samples_b200.ops import native as nv
    from distributed_llm_code_samples_b200.parallel.fused import FOP_SUM, FusedChannel, _owned_cap, _tiles
    from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
    sw = SymmWorld(rank, world, dev)
    M, N, K = 128 * world * 4, 2048, 1024
    worst = 0.0

    def operands(it, r):
        g = torch.Generator(device=dev).manual_seed(1000 * it + r)
        return ((torch.randn((K, M), generator=g, device=dev) * 0.5).to(torch.bfloat16),
                (torch.randn((K, N), generator=g, device=dev) * 0.5).to(torch.bfloat16))

    for owner_mode, bc in ((0, True), (1, False)):
        chan = FusedChannel(sw, _owned_cap([(M, N)], world, owner_mode), _tiles(M, N))
        obuf = sw.alloc(M * N * 2 + 4096)
        out, off = obuf.carve((M, N), torch.bfloat16)
        rows = M if bc else M // world
        row0 = 0 if bc else rank * rows
        tab = obuf.peers_at(off) if bc else (C.c_void_p * world)(*[out.data_ptr() if q == rank else 0 for q in range(world)])
        sw.host_barrier()
        for it in range(3):
            a, b = operands(it, rank)
            out.fill_(float("nan"))
            sw.host_barrier()
            chan.gemm(a, 1, b, 1, M, N, K, tab, N, row0, owner_mode, FOP_SUM, bc)
            torch.cuda.synchronize()
            ref = None
            for q in range(world):
                aq, bq = operands(it, q)
                p = aq.double().t() @ bq.double()
                ref = p if ref is None else ref + p
            got = out[:rows].double()
            want = ref[row0:row0 + rows]
            d = (got - want).abs().reshape(rows // 128, 128, N // 256, 256).amax(dim=(1, 3))
            sc = want.abs().reshape(rows // 128, 128, N // 256, 256).amax(dim=(1, 3)).clamp_min(1e-30)
            worst = max(worst, float((d / sc).max()))
            sw.host_barrier()
    results[rank, 0] = worst
    results[rank, 1] = 1.0 if nv.device_error()[0] == 0 else 0.0
    sw.close()


def test_fused_gradient_sum_matches_fp64_oracle_per_tile():
    """Gradient-level check on real NVLink (the weight-level criterion could hide a dropped tile): the summed
    gradient itself, per tile, within 1e-2 of the tile's max against an fp64 oracle, at every rank."""
    from distributed_llm_code_samples_b200.parallel.launcher import launch
    results = torch.zeros(WORLD, 2).share_memory_()
    launch(WORLD, _grad_worker, (results,), backend="nccl", device="cuda")
    assert (results[:, 1] == 1.0).all(), "device error"
    assert float(results[:, 0].max()) <= 1e-2, results[:, 0].tolist()


def _cfg(**kw):
    from distributed_llm_code_samples_b200.config import TrainConfig
    base = dict(model_size=DM, ffn_dim=4 * DM, layers=3, device="cuda", backend="nccl", world_size=WORLD, dtype="bf16",
                impl="native", lr=0.02)
    base.update(kw)
    return TrainConfig(**base)


def _init():
    from distributed_llm_code_samples_b200.models import ffn
    g = torch.Generator().manual_seed(21)
    lp = ffn.init_stack(3, DM, 4 * DM, g)
    # ONE optimizer step per rank: after a second step the two paths' 1-ulp rounding differences feed back through
    # the gradients (measured: ~1% of weights then differ by up to 4 ulp, uniformly over all tiles, deterministically)
    return [[p.to(torch.bfloat16).float() for p in l] for l in lp], torch.arange(WORLD)


def _rel(a, b):
    worst = 0.0
    for la, lb in zip(a, b):
        for p, q in zip(la, lb):
            worst = max(worst, ((p.float() - q.float()).norm() / (q.float().norm() + 1e-12)).item())
    return worst


def _oracle(lp, seeds, T, method):
    """fp32 CPU reference on the bf16-rounded weights and batches both GPU paths are fed.  DDP/FSDP sum gradients over
    ranks (train_ffns.py:165), which equals ONE step on the concatenation of every rank's batch; MP sees one batch."""
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.ops import get_ops
    from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
    from distributed_llm_code_samples_b200.utils import data
    tr = SingleTrainer(get_ops("torch", "cpu"), TrainConfig(model_size=DM, ffn_dim=4 * DM, layers=3, dtype="fp32", lr=0.02),
                       [[p.clone() for p in l] for l in lp])
    batches = [(x.to(torch.bfloat16).float(), d.to(torch.bfloat16).float()) for x, d in data.mock_data(seeds, T, DM)]
    tr.step(torch.cat([b[0] for b in batches]), torch.cat([b[1] for b in batches]))
    return tr.params


@pytest.mark.parametrize("method", ["ddp", "fsdp", "mp", "mp_sp", "mp_spf"])
def test_fused_trainer_matches_nccl_trainer(method):
    """Same initial weights and data through the fused peer-memory path, the NCCL path and an fp32 oracle.  The two GPU
    paths round differently (fused: fp32 sum of bf16 partials, one rounding; NCCL: bf16 at every hop), and a bf16 weight
    update at this learning rate is itself only good to a few percent, so the criterion is: the fused path is at least
    as close to the oracle as the NCCL path (measured at 4 GPUs: 0.066-0.072 vs 0.073-0.079 relative update error)."""
    from distributed_llm_code_samples_b200.parallel import api
    lp, seeds = _init()
    T = 512
    fn = {"ddp": api.train_ddp, "fsdp": api.train_fsdp, "mp": api.train_tp, "mp_sp": api.train_tp, "mp_spf": api.train_tp}[method]
    # (fsdp: the copy-engine prefetch form here; the in-kernel gather, the default, has its own test below)
    extra = {"mp_sp": {"mp_mode": "sp"}, "mp_spf": {"mp_mode": "spf"}, "fsdp": {"fsdp_gather": "dma"}}.get(method, {})
    if method == "mp_spf":
        T = max(T, 256 * WORLD)    # the all-gather+GEMM kernel needs whole 128-row blocks per token shard (two here)
    if method.startswith("mp"):
        seeds = seeds[:1]          # MP gives every rank ALL the seeds (train_ffns.py:324): keep it to one step as well
    fused = fn(lp, seeds, T, DM, _cfg(comm="fused", **extra))
    nccl = fn(lp, seeds, T, DM, _cfg(comm="nccl", **extra))
    oracle = _oracle(lp, seeds, T, method)
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    uf, un, uo = upd(fused), upd(nccl), upd(oracle)
    assert all(torch.isfinite(u).all() for l in uf for u in l) and all(torch.isfinite(u).all() for l in un for u in l)
    err_f, err_n = _rel(uf, uo), _rel(un, uo)
    assert err_f < 0.35, err_f          # sanity bound only; bf16 weight-grid rounding alone is 5-17 % at this lr
    assert err_f <= 1.3 * err_n + 0.01, (err_f, err_n)
    assert max(u.abs().max().item() for l in uf for u in l) > 0


def test_fsdp_with_in_kernel_forward_gather_matches_nccl():
    """parallel/fused_fsdp_t.py: transposed activations; the weight is the gathered A operand of every forward GEMM
    (row-sharded form) and of the backward GEMMs of the layers the forward pass no longer holds (K-sharded form)."""
    from distributed_llm_code_samples_b200.parallel import api
    lp, seeds = _init()
    T = 512
    fused = api.train_fsdp(lp, seeds, T, DM, _cfg(comm="fused", fsdp_gather="fused"))
    nccl = api.train_fsdp(lp, seeds, T, DM, _cfg(comm="nccl"))
    oracle = _oracle(lp, seeds, T, "fsdp")
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    uf, un, uo = upd(fused), upd(nccl), upd(oracle)
    assert all(torch.isfinite(u).all() for l in uf for u in l)
    err_f, err_n = _rel(uf, uo), _rel(un, uo)
    assert err_f < 0.35 and err_f <= 1.3 * err_n + 0.01, (err_f, err_n)


@pytest.mark.parametrize("mp_mode", ["ar", "sp", "spf"])
def test_fused_mp_with_gelu_and_residual_matches_nccl(mp_mode):
    """The north-star block (Linear -> GELU -> Linear, residual) under fused MP: the residual rides on the kernels'
    in-place update mode (output pre-filled with the residual, alpha = -1)."""
    from distributed_llm_code_samples_b200.parallel import api
    lp, _ = _init()
    seeds = torch.arange(1)
    T = max(512, 256 * WORLD)
    kw = dict(activation="gelu", residual=True, mp_mode=mp_mode, lr=1e-3)
    fused = api.train_tp(lp, seeds, T, DM, _cfg(comm="fused", **kw))
    nccl = api.train_tp(lp, seeds, T, DM, _cfg(comm="nccl", **kw))
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    assert all(torch.isfinite(p).all() for l in fused for p in l)
    assert max(u.abs().max().item() for l in upd(fused) for u in l) > 0
    assert _rel(upd(fused), upd(nccl)) < 0.25


@pytest.mark.parametrize("method", ["ddp", "fsdp", "mp"])
def test_fused_adam_matches_unfused_adam(method):
    """Adam inside the fused kernel's owner reduction (moments sharded by tile owner; MP: fused collectives + the native
    Adam kernel on the shard-local dW) vs the unfused trainer (bf16 all-reduce / reduce-scatter, then b200_adam_bf16).  Adam's first step moves every weight by ~lr regardless of the
    gradient scale, so the two must agree closely; two steps exercise the stored moments."""
    from distributed_llm_code_samples_b200.parallel import api
    lp, _ = _init()
    seeds = torch.arange(2 if method == "mp" else 2 * WORLD)      # two optimizer steps either way (MP: every rank sees every seed)
    fn = {"ddp": api.train_ddp, "fsdp": api.train_fsdp, "mp": api.train_tp}[method]
    fused = fn(lp, seeds, 512, DM, _cfg(comm="fused", optimizer="adam", lr=1e-3))
    unfused = fn(lp, seeds, 512, DM, _cfg(comm="nccl", optimizer="adam", lr=1e-3))
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    assert all(torch.isfinite(p).all() for l in fused for p in l)
    assert max(u.abs().max().item() for l in upd(fused) for u in l) > 0
    assert _rel(upd(fused), upd(unfused)) < 0.25


@pytest.mark.parametrize("method,mp_mode", [("ddp", "ar"), ("fsdp", "ar"), ("mp", "ar"), ("mp", "sp"), ("mp", "spf")])
@pytest.mark.parametrize("optimizer", ["sgd", "adam"])
def test_fused_north_star_block_matches_nccl(method, mp_mode, optimizer):
    """BASELINE.json's block - Linear -> GELU -> Linear + residual, with biases, SGD or Adam - on the FUSED trainers (bias
    in the GEMM epilogues / pre-filled outputs, bias gradients = native column sums + the native fp32 all-reduce) against
    the unfused trainers on NCCL.  Two optimizer steps; weights and biases compared."""
    from distributed_llm_code_samples_b200.models import ffn
    from distributed_llm_code_samples_b200.parallel import api
    g = torch.Generator().manual_seed(44)
    lp = [[p.to(torch.bfloat16).float() if j < 2 else p.normal_(0, 0.05, generator=g) for j, p in enumerate(l)]
          for l in ffn.init_stack(3, DM, 4 * DM, g, bias=True)]
    seeds = torch.arange(2 if method == "mp" else 2 * WORLD)
    T = max(512, 256 * WORLD)
    fn = {"ddp": api.train_ddp, "fsdp": api.train_fsdp, "mp": api.train_tp}[method]
    kw = dict(activation="gelu", residual=True, bias=True, optimizer=optimizer, mp_mode=mp_mode, lr=1e-3)
    fused = fn(lp, seeds, T, DM, _cfg(comm="fused", **kw))
    nccl = fn(lp, seeds, T, DM, _cfg(comm="nccl", **kw))
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    uf, un = upd(fused), upd(nccl)
    assert all(torch.isfinite(p).all() for l in fused for p in l)
    for j in range(4):                                            # W1, W2, b1, b2 separately: a dead bias path must not hide
        a, b = [[l[j]] for l in uf], [[l[j]] for l in un]
        assert max(u[0].abs().max().item() for u in a) > 0, f"param {j} did not move"
        assert _rel(a, b) < 0.25, (j, _rel(a, b))


def test_fused_adam_state_resumes_across_strategies():
    """The fused trainers keep Adam's moments in a compact per-owned-tile layout (1/world per rank); a checkpoint holds
    them in the parameters' full layout.  One fused-DDP step, export, one fused-FSDP step resumed from it, must match two
    uninterrupted steps of the unfused NCCL DDP trainer - and must NOT match a resume that drops the moments."""
    from distributed_llm_code_samples_b200.models import ffn
    from distributed_llm_code_samples_b200.parallel import api
    g = torch.Generator().manual_seed(45)
    lp = [[p.to(torch.bfloat16).float() if j < 2 else p.normal_(0, 0.05, generator=g) for j, p in enumerate(l)]
          for l in ffn.init_stack(3, DM, 4 * DM, g, bias=True)]
    seeds = torch.arange(2 * WORLD)
    kw = dict(optimizer="adam", bias=True, lr=1e-3)
    straight = api.train_ddp(lp, seeds, 512, DM, _cfg(comm="nccl", **kw))
    half = api.train_ddp(lp, seeds[:WORLD], 512, DM, _cfg(comm="fused", **kw))
    st = api.last_optimizer_state()
    assert st["t"] == 1 and all(m.abs().sum() > 0 for l in st["m"] for m in l), "every tile owner must have exported its moments"
    resumed = api.train_fsdp(half, seeds[WORLD:], 512, DM, _cfg(comm="fused", **kw), st)
    cold = api.train_fsdp(half, seeds[WORLD:], 512, DM, _cfg(comm="fused", **kw))
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    assert _rel(upd(resumed), upd(straight)) < 0.2, _rel(upd(resumed), upd(straight))
    assert _rel(upd(cold), upd(straight)) > 1.5 * _rel(upd(resumed), upd(straight))


@pytest.mark.parametrize("method,extra", [("ddp", {"optimizer": "adam"}), ("fsdp", {"bias": True}), ("mp", {"residual": True})])
def test_unfused_trainers_on_native_collectives_match_nccl(method, extra):
    """--comm native: the plain trainers (so Adam / bias / residual work) with every collective on the peer-memory
    kernels instead of NCCL; same weights and data must give the NCCL trainers' result to bf16 accuracy."""
    from distributed_llm_code_samples_b200.models import ffn
    from distributed_llm_code_samples_b200.parallel import api
    g = torch.Generator().manual_seed(33)
    lp = [[p.to(torch.bfloat16).float() for p in l] for l in ffn.init_stack(3, DM, 4 * DM, g, bias=extra.get("bias", False))]
    seeds = torch.arange(1 if method == "mp" else WORLD)
    fn = {"ddp": api.train_ddp, "fsdp": api.train_fsdp, "mp": api.train_tp}[method]
    native = fn(lp, seeds, 512, DM, _cfg(comm="native", lr=1e-3, **extra))
    nccl = fn(lp, seeds, 512, DM, _cfg(comm="nccl", lr=1e-3, **extra))
    upd = lambda res: [[p.float() - q for p, q in zip(a, b)] for a, b in zip(res, lp)]
    assert all(torch.isfinite(p).all() for l in native for p in l)
    assert _rel(upd(native), upd(nccl)) < 0.25


def test_probe_single_process_collectives_and_streams():
    """test_nccl.py / test_torch_cuda_stream.py counterparts: one process, one stream per device."""
    from probes import single_process
    assert all(single_process.collectives(WORLD).values())
    assert all(single_process.streams(WORLD).values())


def test_probe_async_allreduce_and_ipc_barrier():
    """test_torch_distributed.py / test_mp_barrier_gpus.py counterparts: one process per GPU."""
    from probes import multi_process
    assert multi_process.async_allreduce_side_stream(WORLD)
    assert multi_process.ipc_barrier(WORLD)


def _vmm_worker(rank, world, dev, out):
    import ctypes as C
    from distributed_llm_code_samples_b200.ops import native as nv
    from distributed_llm_code_samples_b200.parallel import vmm
    from distributed_llm_code_samples_b200.parallel.symm import CollChannel, SymmWorld
    fd_ok, mc_ok = vmm.supported()
    if not fd_ok:
        out[rank] = -1.0                                   # unsupported on this box: reported as a skip
        return
    n = 1 << 20
    buf = vmm.VmmBuffer(rank, world, dev, n * 2, multicast=mc_ok, tag="t1")
    x = buf.tensor(0, (n,), torch.bfloat16)
    g = torch.Generator().manual_seed(5)
    allx = [torch.randn(n, generator=g).to(torch.bfloat16) for _ in range(world)]
    x.copy_(allx[rank].to(dev))
    torch.cuda.synchronize()
    # (a) the imported peer blocks are usable as an ordinary peer table: native (P2P) all-reduce on them
    sw = SymmWorld(rank, world, dev)
    chan = CollChannel(sw)
    lib = nv.load()
    nv._check(lib.b200_all_reduce_bf16(buf.peers_c, 0, n * 2, chan.flags.peers_c, rank, world, chan.next_epoch(), 0, nv._stream()), "all_reduce on VMM blocks")
    torch.cuda.synchronize()
    total = sum(t.float() for t in allx)
    ok = bool((x.float().cpu() - total).abs().max() <= 0.05 * world)
    # (b) the same sum through the switch
    if mc_ok and buf.mc_ptr:
        x.copy_(allx[rank].to(dev))
        torch.cuda.synchronize()
        nv._check(lib.b200_all_reduce_nvls_bf16(C.c_void_p(buf.mc_ptr), n * 2, chan.flags.peers_c, rank, world, chan.next_epoch(), 0, nv._stream()),
                  "b200_all_reduce_nvls_bf16")
        torch.cuda.synchronize()
        ok &= bool((x.float().cpu() - total).abs().max() <= 0.05 * world)
    ok &= nv.device_error()[0] == 0
    out[rank] = (2.0 if mc_ok else 1.0) if ok else 0.0


def test_vmm_multicast_all_reduce():
    """EXPERIMENTAL runtime (csrc/vmm.cu): cuMem blocks shared by fd, used as a peer table, and - where the box has
    NVLS - an in-switch all-reduce (multimem.ld_reduce / multimem.st) on the multicast mapping."""
    from distributed_llm_code_samples_b200.parallel.launcher import launch
    out = torch.zeros(WORLD).share_memory_()
    launch(WORLD, _vmm_worker, (out,), backend="nccl", device="cuda")
    if (out == -1.0).all():
        pytest.skip("POSIX-fd shareable allocations not supported here")
    assert (out >= 1.0).all(), out.tolist()


def _stress_worker(rank, world, dev, out):
    """Flag-protocol stress: hundreds of back-to-back barriers and small all-reduces on alternating
    sizes with rank-dependent delays in between; any lost or stale flag shows up as a wrong sum or a device timeout."""
    from distributed_llm_code_samples_b200.ops import native as nv
    from distributed_llm_code_samples_b200.parallel.symm import SymmComm, SymmWorld
    sw = SymmWorld(rank, world, dev)
    comm = SymmComm(sw, staging_bytes=4 << 20)
    big = sw.alloc(1 << 20).tensor(0, (1 << 19,), torch.bfloat16)
    ok = True
    junk = torch.empty(1 << 22, device=dev)
    for it in range(300):
        if (it + rank) % 3 == 0:
            junk.normal_()                                  # skew the ranks
        n = (8 * world) << (it % 5)
        t = torch.full((n,), float(rank + 1 + it % 4), dtype=torch.float32, device=dev)
        comm.all_reduce(t)
        if it % 7 == 0:
            comm.barrier()
        big.fill_(1.0)
        comm.all_reduce(big)                                # two-shot path, in-place in symmetric memory
        if it % 50 == 49:
            torch.cuda.synchronize()
            expect = sum(r + 1 + it % 4 for r in range(world))
            ok &= bool((t == expect).all()) and bool((big.float() == world).all())
    torch.cuda.synchronize()
    ok &= nv.device_error()[0] == 0
    out[rank] = 1.0 if ok else 0.0
    sw.close()


def test_barrier_and_flag_protocol_stress():
    from distributed_llm_code_samples_b200.parallel.launcher import launch
    out = torch.zeros(WORLD).share_memory_()
    launch(WORLD, _stress_worker, (out,), backend="nccl", device="cuda")
    assert out.tolist() == [1.0] * WORLD
