"""Fused MP vs NCCL MP vs an fp32 single-device oracle on the same bf16-rounded weights/batches."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models import ffn
from distributed_llm_code_samples_b200.ops import get_ops
from distributed_llm_code_samples_b200.parallel import api
from distributed_llm_code_samples_b200.parallel.single import SingleTrainer
from distributed_llm_code_samples_b200.utils import data

if __name__ == "__main__":
    WORLD = torch.cuda.device_count()
    DM = int(os.environ.get("DM", max(256, 128 * WORLD)))
    T, L, lr = 512, 3, 0.02
    g = torch.Generator().manual_seed(21)
    lp = [[p.to(torch.bfloat16).float() for p in l] for l in ffn.init_stack(L, DM, 4 * DM, g)]
    seeds = torch.arange(1)
    def cfg(comm):
        return TrainConfig(model_size=DM, ffn_dim=4 * DM, layers=L, device="cuda", backend="nccl", world_size=WORLD, dtype="bf16", impl="native", lr=lr, comm=comm)
    fused = api.train_tp(lp, seeds, T, DM, cfg("fused"))
    nccl = api.train_tp(lp, seeds, T, DM, cfg("nccl"))
    # oracle: fp32 math on CPU, bf16-rounded inputs (what both GPU paths are fed)
    ora = SingleTrainer(get_ops("torch", "cpu"), TrainConfig(model_size=DM, ffn_dim=4 * DM, layers=L, dtype="fp32", lr=lr), [[p.clone() for p in l] for l in lp])
    for x, dy in data.mock_data(seeds, T, DM):
        ora.step(x.to(torch.bfloat16).float(), dy.to(torch.bfloat16).float())
    for l in range(L):
        for k in range(2):
            uo = ora.params[l][k] - lp[l][k]
            uf = fused[l][k].float() - lp[l][k]
            un = nccl[l][k].float() - lp[l][k]
            print(json.dumps({"world": WORLD, "DM": DM, "layer": l, "param": k, "rel_fused_vs_oracle": round(((uf - uo).norm() / uo.norm()).item(), 4),
                              "rel_nccl_vs_oracle": round(((un - uo).norm() / uo.norm()).item(), 4), "rel_fused_vs_nccl": round(((uf - un).norm() / un.norm()).item(), 4),
                              "upd_norm": round(uo.norm().item(), 4)}))
