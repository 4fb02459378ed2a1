"""Tiny driver for ncu: runs ONE FFN-shaped GEMM a few times.  usage: python tools/ncu_gemm.py [case] [iters]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.ops import native as nv

case = sys.argv[1] if len(sys.argv) > 1 else "G1"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda:0")
T, D, F = 16384, 4096, 16384
x = torch.randn(T, D, device=dev).to(torch.bfloat16)
w1 = (torch.randn(F, D, device=dev) * 0.013).to(torch.bfloat16)
w2 = (torch.randn(D, F, device=dev) * 0.013).to(torch.bfloat16)
a = torch.randn(T, F, device=dev).to(torch.bfloat16)
dy = torch.randn(T, D, device=dev).to(torch.bfloat16)
h = torch.empty(T, F, dtype=torch.bfloat16, device=dev)
y = torch.empty(T, D, dtype=torch.bfloat16, device=dev)
dw = torch.empty(D, F, dtype=torch.bfloat16, device=dev)
fn = {
    "G1": lambda: nv.gemm(x, 0, w1, 0, h, T, F, D, nv.EPI_RELU),
    "G2": lambda: nv.gemm(a, 0, w2, 0, y, T, D, F),
    "G5": lambda: nv.gemm(dy, 0, w2, 1, h, T, F, D, nv.EPI_RELU_BWD, aux=a),
    "G4": lambda: nv.gemm(dy, 1, a, 1, dw, D, F, T),
    "cublas_G1": lambda: torch.matmul(x, w1.T, out=h),
    "cublas_G2": lambda: torch.matmul(a, w2.T, out=y),
    "cublas_G4": lambda: torch.matmul(dy.T, a, out=dw),
}[case]
for _ in range(iters):
    fn()
torch.cuda.synchronize()
print("ok", case)
