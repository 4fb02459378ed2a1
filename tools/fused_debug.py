"""Compare fused vs NCCL trainers on N GPUs and print WHERE they differ (per layer/param, per 128x256 tile)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.parallel import api

if __name__ == "__main__":
    WORLD = torch.cuda.device_count()
    method = sys.argv[1] if len(sys.argv) > 1 else "ddp"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    DM = int(os.environ.get("DM", max(256, 128 * WORLD)))
    g = torch.Generator().manual_seed(21)
    lp = [[p.to(torch.bfloat16).float() for p in l] for l in ffn.init_stack(3, DM, 4 * DM, g)]
    seeds = torch.arange(steps * WORLD) if method != "mp" else torch.arange(steps)
    T = 512
    def cfg(comm):
        return TrainConfig(model_size=DM, ffn_dim=4 * DM, layers=3, device="cuda", backend="nccl", world_size=WORLD, dtype="bf16", impl="native", lr=0.02, comm=comm)
    fn = {"ddp": api.train_ddp, "fsdp": api.train_fsdp, "mp": api.train_tp}[method]
    fused = fn(lp, seeds, T, DM, cfg("fused"))
    fused2 = fn(lp, seeds, T, DM, cfg("fused"))
    nccl = fn(lp, seeds, T, DM, cfg("nccl"))
    for l in range(3):
        for k in range(2):
            f, f2, n, w0 = fused[l][k].float(), fused2[l][k].float(), nccl[l][k].float(), lp[l][k]
            u = n - w0
            ulp = torch.exp2(torch.floor(torch.log2(n.abs().clamp_min(2.0 ** -20))) - 7)
            ex = ((f - n).abs() - (3 * ulp + 0.05 * u.abs())) > 0
            M, N = f.shape
            tm, tn = (M + 127) // 128, (N + 255) // 256
            tiles = []
            for i in range(tm):
                for j in range(tn):
                    fr = ex[i * 128:(i + 1) * 128, j * 256:(j + 1) * 256].float().mean().item()
                    if fr > 0: tiles.append((i, j, round(fr, 3)))
            print(json.dumps({"nan_fused": int(torch.isnan(f).sum()), "nan_nccl": int(torch.isnan(n).sum()), "layer": l, "param": k, "shape": [M, N], "bad_frac": round(ex.float().mean().item(), 5),
                              "fused_repeatable": bool(torch.equal(f, f2)), "max_abs_diff": (f - n).abs().max().item(),
                              "upd_max": u.abs().max().item(), "bad_tiles": tiles[:12], "n_bad_tiles": len(tiles), "n_tiles": tm * tn}))
