"""Where does the end-to-end loop lose time against the device-timed step?  Under torchrun:

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/e2e_probe.py --method fsdp

Runs the device-timed loop, then Engine.train (public API: pinned-host inputs every step, metric read back every step)
in variants that remove one ingredient at a time, and prints ms/step (max over ranks) for each:
  e2e            the real thing
  no_readback    inputs staged from host, metric computed but never read back (no host sync per step)
  main_stream    host->device copies on the compute stream instead of the prefetcher's side stream
  h2d_only       just the host->device copies of K batches (what PCIe sustains with all ranks copying at once)
"""
import argparse, json, os, sys, time
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as B  # noqa: E402
from distributed_llm_code_samples_b200.config import TrainConfig  # noqa: E402
from distributed_llm_code_samples_b200.engine import Engine, init_params_on_device  # noqa: E402
from distributed_llm_code_samples_b200.utils.data import PinnedPrefetcher  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--method", default="fsdp")
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--mp_input", default="sharded")
ap.add_argument("--fsdp_gather", default="fused")
a = ap.parse_args()
rank, world, local = B.rank_world()
bargs = argparse.Namespace(method=a.method, model_size=None, ffn_dim=None, layers=None)
d, f, l = B.shape_defaults(bargs, a.method)
cfg = TrainConfig(num_steps=a.steps, batch_size=8, seq_len=2048, layers=l, model_size=d, ffn_dim=f, dtype="bf16", device="cuda",
                  backend="nccl", impl="native", mp_input=a.mp_input, fsdp_gather=a.fsdp_gather)
torch.cuda.set_device(local)
eng = Engine(cfg, method=a.method, layers_params=init_params_on_device(cfg, torch.device("cuda", local), seed=1234, scale=B.init_scale(d, f)))
dev = eng.device
pool = B.make_host_pool(3, cfg.tokens, d, 1000 * (rank if a.method in ("ddp", "fsdp") else 0))
dpool = [(x.to(dev), dy.to(dev)) for x, dy in pool]
res = {"method": a.method, "world": world, "mp_input": a.mp_input, "fsdp_gather": a.fsdp_gather}


def batches(n):
    for i in range(n):
        yield pool[i % len(pool)]


def wall(fn):
    B.barrier_sync(world)
    t0 = time.perf_counter()
    fn()
    B.barrier_sync(world)
    return B.max_ranks((time.perf_counter() - t0) * 1e3, world, dev) / a.steps


for i in range(3):
    eng.step(*dpool[i % 3])
res["device_ms"] = wall(lambda: [eng.step(*dpool[i % 3]) for i in range(a.steps)])
for _ in eng.train(batches(2)):
    pass
res["e2e_ms"] = wall(lambda: list(eng.train(batches(a.steps))))


def no_readback():
    for x, dy in PinnedPrefetcher(batches(a.steps), dev):
        eng.ops.sumsq(eng.step(x, dy))


res["no_readback_ms"] = wall(no_readback)


def main_stream():
    for hx, hdy in batches(a.steps):
        y = eng.step(hx.to(dev, non_blocking=True), hdy.to(dev, non_blocking=True))
        float(eng.ops.sumsq(y))


res["main_stream_ms"] = wall(main_stream)


def h2d_only():
    for hx, hdy in batches(a.steps):
        hx.to(dev, non_blocking=True); hdy.to(dev, non_blocking=True)


res["h2d_only_ms"] = wall(h2d_only)
res["e2e_again_ms"] = wall(lambda: list(eng.train(batches(a.steps))))
if rank == 0:
    print(json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in res.items()}), flush=True)
eng.close()
if dist.is_initialized():
    dist.barrier()
    dist.destroy_process_group()
