"""Debug/bench the native collectives on N GPUs: python -m torch.distributed.run --nproc-per-node N tools/coll_debug.py"""
import os, sys, json
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.ops import native as nv
from distributed_llm_code_samples_b200.parallel.launcher import init_from_env
from distributed_llm_code_samples_b200.parallel.symm import SymmComm, SymmWorld

rank, world, dev = init_from_env("cuda", "nccl")
sw = SymmWorld(rank, world, dev)
comm = SymmComm(sw, staging_bytes=64 << 20)
for n in (1 << 16, 1 << 20, (1 << 22) + 16):
    n = n // (8 * world) * (8 * world)
    for rep in range(3):
        g = torch.Generator().manual_seed(100 + n + rep)
        allx = [(torch.randn(n, generator=g)).to(torch.bfloat16) for _ in range(world)]
        total = sum(t.float() for t in allx)
        x = allx[rank].to(dev)
        comm.all_reduce(x)
        torch.cuda.synchronize()
        d = (x.float().cpu() - total).abs()
        bad = (d > 0.1).nonzero().flatten()
        rs = torch.empty(n // world, dtype=torch.bfloat16, device=dev)
        comm.reduce_scatter_into(rs, allx[rank].to(dev))
        torch.cuda.synchronize()
        sl = total[rank * (n // world): (rank + 1) * (n // world)]
        d2 = (rs.float().cpu() - sl).abs()
        bad2 = (d2 > 0.1).nonzero().flatten()
        print(json.dumps({"rank": rank, "n": n, "rep": rep, "ar_max": d.max().item(), "ar_bad": int(bad.numel()),
                          "ar_first": bad[:4].tolist(), "ar_last": bad[-2:].tolist(),
                          "ar_val": [x[i].item() for i in bad[:2].tolist()], "ar_exp": [total[i].item() for i in bad[:2].tolist()],
                          "ar_own": [allx[rank][i].item() for i in bad[:2].tolist()],
                          "rs_max": d2.max().item(), "rs_bad": int(bad2.numel()), "rs_first": bad2[:4].tolist(), "rs_last": bad2[-2:].tolist()}), flush=True)
dist.barrier()
sw.close()
