"""One-shot GEMM diagnosis + timing for a GPU call.

For each operand layout it runs small structured problems whose wrong answers identify WHICH layout assumption is
broken (swizzle, LBO/SBO, K advance, TMEM lane/column mapping), then times the FFN-stack shapes with CUDA events
and prints TFLOP/s next to torch.matmul (cuBLAS) on the same shapes.  Output goes to stdout and, as JSON lines,
to gpurun_out/gemm_diag.jsonl.

    python tools/gemm_diag.py [--no-bench] [--shapes small|ffn]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.ops import native as nv  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")


def run(layout: str, A: torch.Tensor, B: torch.Tensor, M: int, N: int, K: int, **kw) -> torch.Tensor:
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=A.device)
    am = nv.MAJOR_MN if layout in ("NT",) else nv.MAJOR_K
    bm = nv.MAJOR_MN if layout in ("NN", "NT") else nv.MAJOR_K
    nv.gemm(A, am, B, bm, out, M, N, K, **kw)
    torch.cuda.synchronize()
    return out


def operands(layout: str, M: int, N: int, K: int, dev, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(M, K, generator=g)
    b = torch.randn(N, K, generator=g) * 0.25
    a16, b16 = a.to(torch.bfloat16), b.to(torch.bfloat16)
    ref = a16.float() @ b16.float().T
    A = a16.T.contiguous() if layout == "NT" else a16          # MN-major A is stored [K, M]
    Bt = b16.T.contiguous() if layout in ("NN", "NT") else b16  # MN-major B is stored [K, N]
    return A.to(dev), Bt.to(dev), ref.to(dev)


def signature(out: torch.Tensor, ref: torch.Tensor) -> dict:
    """Compact description of where the result is wrong."""
    o = out.float()
    bad = ~torch.isclose(o, ref, rtol=3e-2, atol=3e-2) | ~torch.isfinite(o)
    M, N = bad.shape
    sig = {"bad_frac": round(bad.float().mean().item(), 4), "max_err": float((o - ref).abs().nan_to_num(1e9).max().item())}
    if bad.any():
        rows = bad.any(1).nonzero().flatten()
        cols = bad.any(0).nonzero().flatten()
        sig["bad_rows_first"] = rows[:8].tolist()
        sig["bad_cols_first"] = cols[:8].tolist()
        sig["bad_rows_n"] = int(rows.numel())
        sig["bad_cols_n"] = int(cols.numel())
        sig["nan_frac"] = round((~torch.isfinite(o)).float().mean().item(), 4)
        # per 32-row x 64-col block error map (first 4x4 blocks)
        bm = bad[: min(M, 128), : min(N, 256)].float()
        blocks = bm.reshape(-1, 32, bm.shape[1]).mean(1)
        blocks = blocks.reshape(blocks.shape[0], -1, 64).mean(2)
        sig["block_map_32x64"] = [[round(v, 2) for v in r] for r in blocks.tolist()]
    return sig


def diag(dev) -> bool:
    ok_all = True
    recs = []
    for layout in ("TN", "NN", "NT"):
        for (M, N, K) in [(128, 256, 16), (128, 256, 64), (128, 256, 128), (128, 256, 512), (256, 512, 256), (1024, 1024, 1024)]:
            A, B, ref = operands(layout, M, N, K, dev)
            try:
                out = run(layout, A, B, M, N, K)
                sig = signature(out, ref)
            except Exception as e:  # noqa: BLE001
                sig = {"exception": repr(e)[:300], "device_error": nv.device_error()}
                print(json.dumps({"layout": layout, "shape": [M, N, K], **sig}))
                return False
            ok = sig["bad_frac"] == 0
            ok_all &= ok
            rec = {"layout": layout, "shape": [M, N, K], "ok": ok, **sig}
            recs.append(rec)
            print(json.dumps(rec))
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, "gemm_diag.jsonl"), "w") as f:
        for r in recs:
            f.write(json.dumps(r) + "\n")
    return ok_all


def time_fn(fn, iters=10, warmup=3, flush=None) -> float:
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def bench(dev, which: str) -> None:
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    if which == "small":
        T, D, F = 4096, 2048, 8192
    else:
        T, D, F = 16384, 4096, 16384
    x = torch.randn(T, D, device=dev).to(torch.bfloat16)
    w1 = (torch.randn(F, D, device=dev) * 0.02).to(torch.bfloat16)
    w2 = (torch.randn(D, F, device=dev) * 0.02).to(torch.bfloat16)
    a = torch.randn(T, F, device=dev).to(torch.bfloat16)
    dy = torch.randn(T, D, device=dev).to(torch.bfloat16)
    h = torch.empty(T, F, dtype=torch.bfloat16, device=dev)
    y = torch.empty(T, D, dtype=torch.bfloat16, device=dev)
    dw1 = torch.empty(F, D, dtype=torch.bfloat16, device=dev)
    dw2 = torch.empty(D, F, dtype=torch.bfloat16, device=dev)
    mask = nv.relu_mask_empty(T, F, dev)
    cases = [
        ("G1m h=relu(x@W1^T) TN + sign bitmask (the step's default)", lambda: nv.gemm(x, 0, w1, 0, h, T, F, D, nv.EPI_RELU_MASK, out2=mask), lambda: torch.relu(x @ w1.T), 2 * T * F * D),
        ("G5m da=dy@W2 NN + relu_bwd from bitmask (the step's default)", lambda: nv.gemm(dy, 0, w2, 1, h, T, F, D, nv.EPI_RELU_BWD_MASK, aux=mask), lambda: (dy @ w2).masked_fill_(a <= 0, 0), 2 * T * F * D),
        ("G1 h=relu(x@W1^T) TN", lambda: nv.gemm(x, 0, w1, 0, h, T, F, D, nv.EPI_RELU), lambda: torch.relu(x @ w1.T), 2 * T * F * D),
        ("G2 y=a@W2^T TN", lambda: nv.gemm(a, 0, w2, 0, y, T, D, F), lambda: a @ w2.T, 2 * T * F * D),
        ("G5 da=dy@W2 NN +relu_bwd", lambda: nv.gemm(dy, 0, w2, 1, h, T, F, D, nv.EPI_RELU_BWD, aux=a), lambda: (dy @ w2).masked_fill_(a <= 0, 0), 2 * T * F * D),
        ("G7 dx=dh@W1 NN", lambda: nv.gemm(a, 0, w1, 1, y, T, D, F), lambda: a @ w1, 2 * T * F * D),
        ("G4 dW2=dy^T@a NT", lambda: nv.gemm(dy, 1, a, 1, dw2, D, F, T), lambda: dy.T @ a, 2 * T * F * D),
        ("G6 dW1=dh^T@x NT +sgd", lambda: nv.gemm(a, 1, x, 1, dw1, F, D, T, nv.EPI_SGD, alpha=1e-5), lambda: w1.add_(a.T @ x, alpha=-1e-5), 2 * T * F * D),
    ]
    # fp32 parity mode: split operands, six bf16 plane products (csrc/gemm_f32.cu) vs torch's fp32 GEMM (TF32 off)
    x32, w32 = x.float(), w1.float()
    h32 = torch.empty(T, F, dtype=torch.float32, device=dev)
    X3, W3 = nv.split3(x32), nv.split3(w32)
    torch.backends.cuda.matmul.allow_tf32 = False
    cases.append(("F32 h=relu(x@W1^T) fp32 via 6 bf16 plane products (planes pre-split)", lambda: nv.gemm_f32(X3, 0, W3, 0, h32, T, F, D, nv.EPI_RELU),
                  lambda: torch.relu(x32 @ w32.T), 2 * T * F * D))
    cases.append(("F32 same incl. splitting both operands", lambda: nv.gemm_f32(nv.split3(x32), 0, nv.split3(w32), 0, h32, T, F, D, nv.EPI_RELU),
                  lambda: torch.relu(x32 @ w32.T), 2 * T * F * D))
    recs = []
    for name, mine, ref, flops in cases:
        tm = time_fn(mine, flush=flush)
        tr = time_fn(ref, flush=flush)
        rec = {"case": name, "shape": [T, D, F], "ours_ms": round(tm, 4), "torch_ms": round(tr, 4),
               "ours_tflops": round(flops / tm / 1e9, 1), "torch_tflops": round(flops / tr / 1e9, 1)}
        recs.append(rec)
        print(json.dumps(rec))
    with open(os.path.join(OUT, f"gemm_bench_{which}{os.environ.get('GEMM_BENCH_TAG', '')}.jsonl"), "w") as f:
        for r in recs:
            f.write(json.dumps(r) + "\n")


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-bench", action="store_true")
    ap.add_argument("--shapes", default="ffn")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    print("device:", torch.cuda.get_device_name(0), "sms:", nv.num_sms())
    ok = diag(dev)
    print("DIAG_OK" if ok else "DIAG_FAILED")
    if ok and not args.no_bench:
        bench(dev, args.shapes)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
