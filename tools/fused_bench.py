"""Per-kernel timing of the fused GEMM+collective against its unfused pieces, under torchrun on N GPUs.
Reports max-over-ranks device time (CUDA events), plus the roofline terms: FLOPs at the measured GEMM peak and the
bytes each GPU must send over NVLink at the measured 770 GB/s per direction."""
import ctypes as C, json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.ops import native as nv
from distributed_llm_code_samples_b200.parallel.launcher import init_from_env
from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles

rank, world, dev = init_from_env("cuda", "nccl")
sw = SymmWorld(rank, world, dev)
PEAK_TF, LINK_GBS = 1394.0, 770.0


def timeit(fn, iters=8, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def case_dw(M, N, T, name):
    """dW[M,N] = dy^T[M,T] x[T,N], reduced over ranks, SGD applied."""
    dy = torch.randn(T, M, device=dev).to(torch.bfloat16)
    x = torch.randn(T, N, device=dev).to(torch.bfloat16)
    wb = sw.alloc(M * N * 2 + 4096)
    W, off = wb.carve((M, N), torch.bfloat16)
    W.normal_(0, 0.02)
    tab = wb.peers_at(off)
    g = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    ch0 = FusedChannel(sw, _owned_cap([(M, N)], world, 0), _tiles(M, N))
    res = {"case": name, "world": world, "M": M, "N": N, "K": T}
    res["local_sgd_gemm_ms"] = timeit(lambda: nv.gemm(dy, 1, x, 1, W, M, N, T, nv.EPI_SGD, alpha=1e-6))
    res["fused_allreduce_sgd_ms"] = timeit(lambda: ch0.gemm(dy, 1, x, 1, M, N, T, tab, N, 0, 0, 1, True, 1e-6))
    if M % (128 * world) == 0:
        ch1 = FusedChannel(sw, _owned_cap([(M, N)], world, 1), _tiles(M, N))
        Ws = W[rank * (M // world):(rank + 1) * (M // world)]
        ltab = (C.c_void_p * world)(*[Ws.data_ptr() if r == rank else 0 for r in range(world)])
        res["fused_reducescatter_sgd_ms"] = timeit(lambda: ch1.gemm(dy, 1, x, 1, M, N, T, ltab, N, rank * (M // world), 1, 1, False, 1e-6))

    def unfused():
        nv.gemm(dy, 1, x, 1, g, M, N, T)
        dist.all_reduce(g)
        nv.sgd_(W, g, 1e-6)
    res["gemm_nccl_allreduce_sgd_ms"] = timeit(unfused)
    res["nccl_allreduce_only_ms"] = timeit(lambda: dist.all_reduce(g))
    flops = 2.0 * M * N * T
    res["compute_floor_ms"] = flops / (PEAK_TF * 1e12) * 1e3
    res["link_floor_allreduce_ms"] = 2.0 * (world - 1) / world * M * N * 2 / (LINK_GBS * 1e9) * 1e3
    res["roofline_frac_fused_allreduce"] = max(res["compute_floor_ms"], res["link_floor_allreduce_ms"]) / res["fused_allreduce_sgd_ms"]
    return res


def case_mp(T, D, Fs, name):
    """y[T,D] = a[T,Fs] W2s[D,Fs]^T summed over ranks (row-parallel GEMM + all-reduce)."""
    a = torch.randn(T, Fs, device=dev).to(torch.bfloat16)
    w = (torch.randn(D, Fs, device=dev) * 0.02).to(torch.bfloat16)
    yb = sw.alloc(T * D * 2 + 4096)
    y, off = yb.carve((T, D), torch.bfloat16)
    tab = yb.peers_at(off)
    ch = FusedChannel(sw, _owned_cap([(T, D)], world, 0), _tiles(T, D))
    y2 = torch.empty(T, D, dtype=torch.bfloat16, device=dev)
    res = {"case": name, "world": world, "M": T, "N": D, "K": Fs}
    res["local_gemm_ms"] = timeit(lambda: nv.gemm(a, 0, w, 0, y2, T, D, Fs))
    res["fused_gemm_allreduce_ms"] = timeit(lambda: ch.gemm(a, 0, w, 0, T, D, Fs, tab, D, 0, 0, 0, True))
    if sw.multicast_supported():           # the same call with the partial tiles summed inside the NVSwitch
        ym = sw.alloc(T * D * 2 + 4096, multicast=True)
        _, offm = ym.carve((T, D), torch.bfloat16)
        chn = FusedChannel(sw, _owned_cap([(T, D)], world, 0), _tiles(T, D), nvls=True)
        res["fused_gemm_allreduce_nvls_ms"] = timeit(lambda: chn.gemm(a, 0, w, 0, T, D, Fs, ym.peers_at(offm), D, 0, 0, 0, True, out_mc=ym.mc_at(offm)))

    def unfused():
        nv.gemm(a, 0, w, 0, y2, T, D, Fs)
        dist.all_reduce(y2)
    res["gemm_nccl_allreduce_ms"] = timeit(unfused)
    res["nccl_allreduce_only_ms"] = timeit(lambda: dist.all_reduce(y2))
    res["compute_floor_ms"] = 2.0 * T * D * Fs / (PEAK_TF * 1e12) * 1e3
    res["link_floor_allreduce_ms"] = 2.0 * (world - 1) / world * T * D * 2 / (LINK_GBS * 1e9) * 1e3
    res["roofline_frac_fused"] = max(res["compute_floor_ms"], res["link_floor_allreduce_ms"]) / res["fused_gemm_allreduce_ms"]
    if "fused_gemm_allreduce_nvls_ms" in res:
        # in-switch reduction: ~ (world-1)/world * S out (the pulls) + S/world (the multicast store) per GPU and direction
        res["link_floor_nvls_ms"] = T * D * 2 / (LINK_GBS * 1e9) * 1e3
        res["roofline_frac_fused_nvls"] = max(res["compute_floor_ms"], res["link_floor_nvls_ms"]) / res["fused_gemm_allreduce_nvls_ms"]
    return res


def case_ag(T, D, Fs, name):
    """a[T,Fs] = relu(all_gather(x shards)[T,D] W1s[Fs,D]^T): all-gather + column-parallel GEMM (csrc/gather_gemm.cu), and the
    reduce-scatter form of the row-parallel GEMM that feeds it.  FUSED_BENCH_ONLY=ag (not in the default run: the kernel
    had not run on hardware when this case was written)."""
    from distributed_llm_code_samples_b200.parallel.fused import GatherChannel
    from distributed_llm_code_samples_b200.parallel.symm import CollChannel
    xb = sw.alloc(T * D * 2 + 4096)
    x, off = xb.carve((T, D), torch.bfloat16)
    x.normal_(0, 1)
    w1 = (torch.randn(Fs, D, device=dev) * 0.02).to(torch.bfloat16)
    a = torch.empty(T, Fs, dtype=torch.bfloat16, device=dev)
    mask = nv.relu_mask_empty(T, Fs, dev)
    gch = GatherChannel(sw, T)
    coll = CollChannel(sw)
    peers = xb.peers_at(off)
    Ts = T // world
    res = {"case": name, "world": world, "M": T, "N": Fs, "K": D}
    res["local_gemm_ms"] = timeit(lambda: nv.gemm(x, 0, w1, 0, a, T, Fs, D, nv.EPI_RELU_MASK, out2=mask))
    res["fused_allgather_gemm_ms"] = timeit(lambda: gch.gemm(peers, w1, 0, a, T, Fs, D, nv.EPI_RELU_MASK, out2=mask))

    def native_unfused():
        nv._check(nv.load().b200_all_gather(xb.peers_c, off, T * D * 2, coll.flags.peers_c, rank, world, coll.next_epoch(), 0, nv._stream()), "ag")
        nv.gemm(x, 0, w1, 0, a, T, Fs, D, nv.EPI_RELU_MASK, out2=mask)
    res["native_allgather_then_gemm_ms"] = timeit(native_unfused)
    full = torch.empty(T, D, dtype=torch.bfloat16, device=dev)
    shard = x[rank * Ts:(rank + 1) * Ts]

    def nccl_unfused():
        dist.all_gather_into_tensor(full, shard)
        nv.gemm(full, 0, w1, 0, a, T, Fs, D, nv.EPI_RELU_MASK, out2=mask)
    res["nccl_allgather_then_gemm_ms"] = timeit(nccl_unfused)
    res["compute_floor_ms"] = 2.0 * T * D * Fs / (PEAK_TF * 1e12) * 1e3
    res["link_floor_allgather_ms"] = (world - 1) / world * T * D * 2 / (LINK_GBS * 1e9) * 1e3
    res["roofline_frac_fused"] = max(res["compute_floor_ms"], res["link_floor_allgather_ms"]) / res["fused_allgather_gemm_ms"]
    return res


out = []
only = os.environ.get("FUSED_BENCH_ONLY", "")
if only == "ag":
    out.append(case_ag(16384, 8192, 28672 // world, "mp-sp a = relu(AG(x) W1^T) (70B shape)"))
if only in ("", "dw"):
    out.append(case_dw(16384, 4096, 16384, "ddp/fsdp dW1 [F,D]"))
    out.append(case_dw(4096, 16384, 16384, "ddp/fsdp dW2 [D,F]"))
if only in ("", "mp"):
    out.append(case_mp(16384, 8192, 28672 // world, "mp y = a W2^T (70B shape)"))
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    tag = os.environ.get("FUSED_BENCH_TAG", "")
    with open(f"gpurun_out/fused_bench_n{world}{tag}.jsonl", "w") as f:
        for r in out:
            r = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}
            print(json.dumps(r)); f.write(json.dumps(r) + "\n")
dist.barrier()
sw.close()
