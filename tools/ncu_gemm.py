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
    "G1m": None, "G5m": None,
    "G7": lambda: nv.gemm(a, 0, w1, 1, y, T, D, F),                                  # NN: B operand MN-major
    "G6": lambda: nv.gemm(a, 1, x, 1, w1, F, D, T, nv.EPI_SGD, alpha=1e-6),          # NT + SGD epilogue (K = tokens)
    "fused_mp_1cta": None, "emu8_dw": None, "emu8_mp": None, "philox": None, "colsum": None, "f32": None,
    "G4": lambda: nv.gemm(dy, 1, a, 1, dw, D, F, T),
    "cublas_G1": lambda: torch.matmul(x, w1.T, out=h),
    "cublas_G2": lambda: torch.matmul(a, w2.T, out=y),
    "cublas_G4": lambda: torch.matmul(dy.T, a, out=dw),
    "fused_dw": None,
    "fused_adam": None,
    "gather": None,
    "sgd": None,
    "adam": None,
    "sumsq": None,
    "coll_ar": None,
    "coll_ag": None,
}[case]
if case in ("G1m", "G5m"):
    mask = nv.relu_mask_empty(T, F, dev)
    nv.gemm(x, 0, w1, 0, h, T, F, D, nv.EPI_RELU_MASK, out2=mask)
    fn = (lambda: nv.gemm(x, 0, w1, 0, h, T, F, D, nv.EPI_RELU_MASK, out2=mask)) if case == "G1m" else \
         (lambda: nv.gemm(dy, 0, w2, 1, h, T, F, D, nv.EPI_RELU_BWD_MASK, aux=mask))
if case == "fused_mp_1cta":
    # the double-buffered 1-CTA fused kernel MP uses at 8 GPUs (70B shard shape), in a world of one
    from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
    from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles
    nv.load().b200_fused_config(4, 1)
    Tm, Dm, Fs = 16384, 8192, 3584
    am = torch.randn(Tm, Fs, device=dev).to(torch.bfloat16)
    wm = (torch.randn(Dm, Fs, device=dev) * 0.02).to(torch.bfloat16)
    sw = SymmWorld(0, 1, dev)
    chan = FusedChannel(sw, _owned_cap([(Tm, Dm)], 1, 0), _tiles(Tm, Dm))
    yb = sw.alloc(Tm * Dm * 2 + 4096)
    _, off = yb.carve((Tm, Dm), torch.bfloat16)
    tab = yb.peers_at(off)
    fn = lambda: chan.gemm(am, 0, wm, 0, Tm, Dm, Fs, tab, Dm, 0, 0, 0, True)
if case in ("emu8_dw", "emu8_mp"):
    # EIGHT emulated ranks in one cooperative launch (parallel/emu.py): the real 8-rank protocol - owner maps, inbox slots
    # per source, flags, broadcast of reduced tiles - with local memory in place of NVLink; 18 CTAs per rank
    from distributed_llm_code_samples_b200.parallel.emu import EmuFused
    lib = nv.load()
    W = 8
    if case == "emu8_dw":
        M_, N_, K_, am_, bm_ = 4096, 4096, 4096, 1, 1
        A = [torch.randn(K_, M_, device=dev).to(torch.bfloat16) for _ in range(W)]
        B = [torch.randn(K_, N_, device=dev).to(torch.bfloat16) for _ in range(W)]
        op = 1
    else:
        M_, N_, K_, am_, bm_ = 4096, 4096, 3584, 0, 0
        A = [torch.randn(M_, K_, device=dev).to(torch.bfloat16) for _ in range(W)]
        B = [(torch.randn(N_, K_, device=dev) * 0.02).to(torch.bfloat16) for _ in range(W)]
        op = 0
        lib.b200_fused_config(4, 1)
    tiles = (M_ // 128) * (N_ // 256)
    em = EmuFused(W, int(lib.b200_fused_owned_tiles(M_, N_, W, 0)), tiles, dev)
    out = [torch.zeros(M_, N_, dtype=torch.bfloat16, device=dev) for _ in range(W)]
    fn = lambda: em.gemm(A, am_, B, bm_, M_, N_, K_, out, 0, op, True, alpha=1e-6)
if case == "f32":
    X3, W3 = nv.split3(x.float()), nv.split3(w1.float())
    h32 = torch.empty(T, F, dtype=torch.float32, device=dev)
    fn = lambda: nv.gemm_f32(X3, 0, W3, 0, h32, T, F, D, nv.EPI_RELU)
if case == "philox":
    buf = torch.empty(T * D, dtype=torch.bfloat16, device=dev)
    fn = lambda: nv.randn_(buf, 1234, 1.0)
if case == "colsum":
    fn = lambda: nv.colsum(a)
if case == "fused_dw":
    # the fused GEMM + collective kernel in a world of ONE: pushes to its own inbox, side warps reduce + apply SGD
    import ctypes as C
    from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
    from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles
    sw = SymmWorld(0, 1, dev)
    chan = FusedChannel(sw, _owned_cap([(D, F)], 1, 0), _tiles(D, F))
    wbuf = sw.alloc(D * F * 2 + 4096)
    W, off = wbuf.carve((D, F), torch.bfloat16)
    tab = wbuf.peers_at(off)
    fn = lambda: chan.gemm(dy, 1, a, 1, D, F, T, tab, F, 0, 0, 1, True, 1e-6)
if case == "fused_adam":
    # same kernel family with the Adam update in the owner's reduction (world of one)
    from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
    from distributed_llm_code_samples_b200.parallel.fused import FusedChannel, _owned_cap, _tiles
    sw = SymmWorld(0, 1, dev)
    chan = FusedChannel(sw, _owned_cap([(D, F)], 1, 0), _tiles(D, F))
    wbuf = sw.alloc(D * F * 2 + 4096)
    W, off = wbuf.carve((D, F), torch.bfloat16)
    tab = wbuf.peers_at(off)
    state = chan.adam_state(D, F, 0)
    fn = lambda: chan.gemm_adam(dy, 1, a, 1, D, F, T, tab, F, 0, 0, True, state, 1e-6, 1)
if case == "gather":
    # the all-gather + GEMM kernel in a world of ONE: no remote rows, so this shows the cost of its extra warps and of the
    # producer's hook on the plain G1 problem (compare with case G1)
    from distributed_llm_code_samples_b200.parallel.symm import SymmWorld
    from distributed_llm_code_samples_b200.parallel.fused import GatherChannel
    sw = SymmWorld(0, 1, dev)
    gch = GatherChannel(sw, T)
    xb = sw.alloc(T * D * 2 + 4096)
    xs, off = xb.carve((T, D), torch.bfloat16)
    xs.copy_(x)
    peers = xb.peers_at(off)
    fn = lambda: gch.gemm(peers, w1, 0, h, T, F, D, nv.EPI_RELU)
if case in ("sgd", "adam", "sumsq"):
    # the standalone elementwise kernels at one weight's size (64 Mi elements)
    g32 = torch.randn(D, F, device=dev)
    m, v = torch.zeros(D, F, device=dev), torch.zeros(D, F, device=dev)
    fn = {"sgd": lambda: nv.sgd_(w2, g32, 1e-6), "adam": lambda: nv.adam_(w2, g32, m, v, 1e-6, 0.9, 0.999, 1e-8, 1),
          "sumsq": lambda: nv.sumsq(w2)}[case]
if case in ("coll_ar", "coll_ag"):
    # the standalone collective kernels in a world of ONE (flag protocol against itself, local loads instead of peer loads):
    # shows their instruction mix and HBM-side behaviour; NVLink-side numbers come from tools/coll_sweep.py
    from distributed_llm_code_samples_b200.parallel.symm import SymmComm, SymmWorld
    sw = SymmWorld(0, 1, dev)
    comm = SymmComm(sw, staging_bytes=1 << 20)
    buf = sw.alloc(128 << 20)
    t = buf.tensor(0, (64 << 20,), torch.bfloat16)
    full = torch.empty(64 << 20, dtype=torch.bfloat16, device=dev)
    fn = (lambda: comm.all_reduce(t)) if case == "coll_ar" else (lambda: comm.all_gather_into(full, t))
for _ in range(iters):
    fn()
torch.cuda.synchronize()
print("ok", case)
