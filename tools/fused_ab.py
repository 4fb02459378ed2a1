"""Single-GPU A/B of the fused dW kernel's overhead against the local SGD GEMM (the 1 -> 2 GPU efficiency loss of round 1
was this kernel's own cost, not NVLink: profiles/ncu/prof_fused_dw_world1.summary.csv).

    python tools/fused_ab.py [tag]        # one JSON line per (shape, variant); B200_NATIVE_LIB selects an A/B build

Variants: side warps 4 (448 threads, 128 registers) / 2 (384 threads, 168 registers), in a world of one (the rank pushes
to its own inbox: the whole push -> flag -> reduce -> SGD -> store path with no link) and as 2 emulated ranks on 74 SMs
each (parallel/emu.py; compare with the local kernel capped at 74 CTAs)."""
import ctypes as C, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.ops import native as nv
from distributed_llm_code_samples_b200.parallel.emu import EmuFused
from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles
from distributed_llm_code_samples_b200.parallel.symm import SymmWorld

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
sw = SymmWorld(0, 1, dev)
lib = nv.load()
tag = sys.argv[1] if len(sys.argv) > 1 else "base"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        nv.fill_zero_(flush)                                   # L2 flush between timed iterations
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def run(M, N, T, name):
    dy = torch.randn(T, M, device=dev).to(torch.bfloat16)
    x = torch.randn(T, N, device=dev).to(torch.bfloat16)
    W = (torch.randn(M, N, device=dev) * 0.02).to(torch.bfloat16)
    tab = (C.c_void_p * 1)(W.data_ptr())
    res = {"tag": tag, "case": name, "M": M, "N": N, "K": T}
    res["local_sgd_ms"] = timeit(lambda: nv.gemm(dy, 1, x, 1, W, M, N, T, nv.EPI_SGD, alpha=1e-6))
    res["local_sgd_74cta_ms"] = timeit(lambda: nv.gemm(dy, 1, x, 1, W, M, N, T, nv.EPI_SGD, alpha=1e-6, max_ctas=74))
    for side in (4, 2):
        lib.b200_fused_config(side, 0)
        ch = FusedChannel(sw, _owned_cap([(M, N)], 1, 0), _tiles(M, N))
        res[f"fused_world1_side{side}_ms"] = timeit(lambda: ch.gemm(dy, 1, x, 1, M, N, T, tab, N, 0, 0, 1, True, 1e-6))
        del ch
        # two emulated ranks, 74 CTAs each, both computing this shape: compare with local_sgd_74cta_ms
        em = EmuFused(2, int(lib.b200_fused_owned_tiles(M, N, 2, 0)), _tiles(M, N), dev)
        W2 = [W, W.clone()]
        res[f"fused_emu2_side{side}_ms"] = timeit(lambda: em.gemm([dy, dy], 1, [x, x], 1, M, N, T, W2, 0, 1, True, alpha=1e-6), iters=6)
        del em, W2
    lib.b200_fused_config(4, -1)
    return res


os.makedirs("gpurun_out", exist_ok=True)
with open(f"gpurun_out/fused_ab_{tag}.jsonl", "w") as f:
    for case in [(16384, 4096, 16384, "dW1 [F,D]"), (4096, 16384, 16384, "dW2 [D,F]")]:
        r = run(*case)
        r = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}
        print(json.dumps(r)); f.write(json.dumps(r) + "\n")
assert nv.device_error()[0] == 0, nv.device_error()
