"""Steady-state (power-capped) throughput of each FFN GEMM: run every case back to back for ~1.5 s and time the last
second with CUDA events.  Prints ours vs torch.matmul (cuBLAS) TFLOP/s and the SM clock seen during the window."""
import json, os, subprocess, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.ops import native as nv

dev = torch.device("cuda:0")
T, D, F = 16384, 4096, 16384
x = torch.randn(T, D, device=dev).to(torch.bfloat16)
w1 = (torch.randn(F, D, device=dev) * 0.013).to(torch.bfloat16)
w2 = (torch.randn(D, F, device=dev) * 0.013).to(torch.bfloat16)
a = torch.relu(torch.randn(T, F, device=dev)).to(torch.bfloat16)
dy = torch.randn(T, D, device=dev).to(torch.bfloat16)
h = torch.empty(T, F, dtype=torch.bfloat16, device=dev)
y = torch.empty(T, D, dtype=torch.bfloat16, device=dev)
dw1 = torch.empty(F, D, dtype=torch.bfloat16, device=dev)
dw2 = torch.empty(D, F, dtype=torch.bfloat16, device=dev)
FL = 2 * T * F * D
cases = {
    "G1": (lambda: nv.gemm(x, 0, w1, 0, h, T, F, D, nv.EPI_RELU), lambda: torch.relu(x @ w1.T)),
    "G1_store": (lambda: nv.gemm(x, 0, w1, 0, h, T, F, D), lambda: torch.matmul(x, w1.T, out=h)),
    "G2": (lambda: nv.gemm(a, 0, w2, 0, y, T, D, F), lambda: torch.matmul(a, w2.T, out=y)),
    "G5": (lambda: nv.gemm(dy, 0, w2, 1, h, T, F, D, nv.EPI_RELU_BWD, aux=a), lambda: torch.matmul(dy, w2, out=h)),
    "G5_store": (lambda: nv.gemm(dy, 0, w2, 1, h, T, F, D), lambda: torch.matmul(dy, w2, out=h)),
    "G7": (lambda: nv.gemm(a, 0, w1, 1, y, T, D, F), lambda: torch.matmul(a, w1, out=y)),
    "G4": (lambda: nv.gemm(dy, 1, a, 1, dw2, D, F, T), lambda: torch.matmul(dy.T, a, out=dw2)),
    "G6_sgd": (lambda: nv.gemm(a, 1, x, 1, dw1, F, D, T, nv.EPI_SGD, alpha=1e-5), lambda: torch.matmul(a.T, x, out=dw1)),
}
want = sys.argv[1].split(",") if len(sys.argv) > 1 else list(cases)
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0


def clock():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, timeout=5).stdout
        c, p = out.strip().split(",")
        return float(c), float(p)
    except Exception:
        return None, None


def sustained(fn):
    t0 = time.time()
    while time.time() - t0 < 0.6 * secs:      # heat up to the power cap
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    s.record()
    t0 = time.time()
    ck = None
    while time.time() - t0 < secs:
        for _ in range(20):
            fn()
        n += 20
        if ck is None and time.time() - t0 > 0.5 * secs:
            ck = clock()
        torch.cuda.synchronize()
    e.record()
    torch.cuda.synchronize()
    return FL * n / (s.elapsed_time(e) * 1e-3) / 1e12, ck


for name in want:
    mine, ref = cases[name]
    tm, cm = sustained(mine)
    tr, cr = sustained(ref)
    print(json.dumps({"case": name, "ours_tflops": round(tm, 1), "ours_clock_power": cm, "cublas_tflops": round(tr, 1), "cublas_clock_power": cr,
                      "pair": os.environ.get("B200_GEMM_PAIR", "1")}), flush=True)
