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
sym = sw.alloc(16 << 20)
n = 1 << 20
for mode in ("staged", "staged_sync", "direct", "direct_hostbarrier"):
    for rep in range(2):
        g = torch.Generator().manual_seed(7 + rep)
        allx = [(torch.randn(n, generator=g)).to(torch.bfloat16) for _ in range(world)]
        total = sum(t.float() for t in allx)
        if mode.startswith("staged"):
            x = allx[rank].to(dev)
            if mode == "staged_sync":
                torch.cuda.synchronize(); dist.barrier()
            comm.all_reduce(x)
            torch.cuda.synchronize()
        else:
            x = sym.tensor(0, (n,), torch.bfloat16)
            x.copy_(allx[rank].to(dev))
            torch.cuda.synchronize()
            if mode == "direct_hostbarrier":
                dist.barrier()
            comm.all_reduce(x)
            torch.cuda.synchronize()
        res = x.float().cpu()
        d = (res - total).abs()
        bad = (d > 0.1).nonzero().flatten()
        own = allx[rank].float(); peer = allx[1 - rank].float()
        is_own = int(((res - own).abs() < 1e-6)[bad].sum()) if bad.numel() else 0
        is_peer = int(((res - peer).abs() < 1e-6)[bad].sum()) if bad.numel() else 0
        segs = []
        if bad.numel():
            b = bad.tolist(); start = b[0]; prev = b[0]
            for v in b[1:]:
                if v > prev + 64:
                    segs.append((start, prev)); start = v
                prev = v
            segs.append((start, prev))
        print(json.dumps({"rank": rank, "mode": mode, "rep": rep, "bad": int(bad.numel()), "is_own": is_own, "is_peer": is_peer, "segs": segs[:6], "nsegs": len(segs)}), flush=True)
# stream-order probe: long native kernel followed at once by a torch op on the same stream
M = 8192
a = torch.randn(M, M, device=dev).to(torch.bfloat16); b = torch.randn(M, M, device=dev).to(torch.bfloat16)
out = torch.zeros(M, M, dtype=torch.bfloat16, device=dev)
torch.cuda.synchronize()
for rep in range(3):
    out.zero_()
    nv.gemm(a, 0, b, 0, out, M, M, M)
    snap = out.clone()
    torch.cuda.synchronize()
    print(json.dumps({"rank": rank, "mode": "order_probe", "rep": rep, "snapshot_equal": bool(torch.equal(snap, out)), "zeros_in_snapshot": int((snap == 0).sum())}), flush=True)
dist.barrier()
sw.close()
