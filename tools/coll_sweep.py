"""Collective sweep (BASELINE.json config 5): native peer-memory all-reduce / all-gather / reduce-scatter vs NCCL via
torch.distributed, 1 KB - 1 GB, device-timed, max over ranks.  Run under torchrun on 2/4/8 GPUs."""
import json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_code_samples_b200.parallel.launcher import init_from_env
from distributed_llm_code_samples_b200.parallel.symm import SymmComm, SymmWorld

rank, world, dev = init_from_env("cuda", "nccl")
sw = SymmWorld(rank, world, dev)
comm = SymmComm(sw, staging_bytes=1 << 20)
max_bytes = int(sys.argv[1]) if len(sys.argv) > 1 else (1 << 30)
buf = sw.alloc(max_bytes + 4096, multicast=True)        # NVLS where the box has it: all-reduce / reduce-scatter sum in the switch
p2p = sw.alloc(max_bytes + 4096)                        # plain IPC block: the peer-pointer kernels, for comparison


def timeit(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


rows = []
size = 1024
while size <= max_bytes:
    n = size // 2
    n = max(n // (8 * world) * (8 * world), 8 * world)
    nbytes = n * 2
    x = buf.tensor(0, (n,), torch.bfloat16)            # symmetric: the native kernels work in place
    x.normal_()
    y = torch.randn(n, device=dev).to(torch.bfloat16)
    shard = torch.empty(n // world, dtype=torch.bfloat16, device=dev)
    full = torch.empty(n, dtype=torch.bfloat16, device=dev)
    iters = 50 if nbytes <= (1 << 20) else (10 if nbytes <= (1 << 26) else 4)
    r = {"bytes": nbytes, "world": world}
    r["nvls"] = bool(getattr(buf, "mc_ptr", None))
    r["ar_native_us"] = timeit(lambda: comm.all_reduce(x), iters) * 1e3
    xp = p2p.tensor(0, (n,), torch.bfloat16)
    xp.normal_()
    r["ar_p2p_us"] = timeit(lambda: comm.all_reduce(xp), iters) * 1e3
    r["ar_nccl_us"] = timeit(lambda: dist.all_reduce(y), iters) * 1e3
    xs = x[rank * (n // world):(rank + 1) * (n // world)]
    r["ag_native_us"] = timeit(lambda: comm.all_gather_into(full, xs), iters) * 1e3
    r["ag_nccl_us"] = timeit(lambda: dist.all_gather_into_tensor(full, shard), iters) * 1e3
    r["rs_native_us"] = timeit(lambda: comm.reduce_scatter_into(shard, x), iters) * 1e3
    r["rs_nccl_us"] = timeit(lambda: dist.reduce_scatter_tensor(shard, y), iters) * 1e3
    bus = 2.0 * (world - 1) / world * nbytes
    r["ar_native_busbw_GBs"] = bus / (r["ar_native_us"] * 1e-6) / 1e9
    r["ar_nccl_busbw_GBs"] = bus / (r["ar_nccl_us"] * 1e-6) / 1e9
    rows.append({k: (round(v, 2) if isinstance(v, float) else v) for k, v in r.items()})
    if rank == 0:
        print(json.dumps(rows[-1]), flush=True)
    size *= 4
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/coll_sweep_n{world}.jsonl", "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
dist.barrier()
sw.close()
