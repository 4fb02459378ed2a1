"""Is the fp32 (split-operand) training step deterministic, and how far is it from an fp64 run?  1 GPU."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import get_ops, native as nv
from distributed_llm_code_samples_b200.parallel.base import place
from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
dev = torch.device("cuda", 0)
torch.backends.cuda.matmul.allow_tf32 = False
g = torch.Generator().manual_seed(3)
lp = ffn.init_stack(3, 256, 1024, g)
x, dy = torch.randn(512, 256, generator=g).to(dev), (0.1 * torch.randn(512, 256, generator=g)).to(dev)
cfg = TrainConfig(model_size=256, ffn_dim=1024, dtype="fp32", lr=0.05)
a, b = torch.randn(512, 256, device=dev), torch.randn(1024, 256, device=dev)
o1, o2 = torch.empty(512, 1024, device=dev), torch.empty(512, 1024, device=dev)
nv.gemm_f32(nv.split3(a), 0, nv.split3(b), 0, o1, 512, 1024, 256)
nv.gemm_f32(nv.split3(a), 0, nv.split3(b), 0, o2, 512, 1024, 256)
print("gemm_f32 bitwise repeatable:", torch.equal(o1, o2), "err vs fp64:", ((o1.double() - a.double() @ b.double().T).abs().max() / (a.double().abs() @ b.double().abs().T).max()).item())
runs = []
for rep in range(3):
    nat = SingleTrainer(get_ops("native", dev), cfg, place(lp, dev, torch.float32))
    o64 = SingleTrainer(get_ops("torch", dev), cfg, [[p.to(dev, torch.float64) for p in l] for l in lp])
    for s in range(2):
        y = nat.step(x, dy); y64 = o64.step(x.double(), dy.double())
        errs = [((p.double() - r).abs().max() / r.abs().max()).item() for A, C in zip(nat.params, o64.params) for p, r in zip(A, C)]
        print(f"rep {rep} step {s}: y err {((y.double() - y64).abs().max() / y64.abs().max()).item():.2e}  weights", " ".join(f"{e:.1e}" for e in errs))
    runs.append([p.clone() for A in nat.params for p in A])
print("training run bitwise repeatable:", all(torch.equal(p, q) for p, q in zip(runs[0], runs[1])) and all(torch.equal(p, q) for p, q in zip(runs[0], runs[2])))
# how many ReLU decisions differ from the fp64 run in the first layer's forward?
h32 = torch.empty(512, 1024, device=dev)
W1 = lp[0][0].to(dev)
nv.gemm_f32(nv.split3(x), 0, nv.split3(W1), 0, h32, 512, 1024, 256)
h64 = x.double() @ W1.double().T
print("sign flips in layer-0 pre-activation vs fp64:", int(((h32 > 0) != (h64 > 0)).sum()), "of", h32.numel(), " cuBLAS fp32:", int((((x @ W1.T) > 0) != (h64 > 0)).sum()))
