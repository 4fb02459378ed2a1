#!/usr/bin/env python
"""Headline benchmark: FFN-stack training tokens/sec (BASELINE.json), device-timed, max over ranks.

    python bench.py --gpus N --steps K --warmup W                  # this framework
    python bench.py --impl reference --gpus N --steps K --warmup W   # the unmodified reference (baseline/_ref)

For N > 1 launch one rank per GPU with torchrun (``python -m torch.distributed.run --nproc-per-node N ...``).
Default config = BASELINE.json's DDP config: bf16, d_model 4096, d_ff 16384, 32 blocks, 8 x 2048 tokens per rank
(weak scaling: per-GPU work fixed).  ``--method fsdp|mp|single`` and the shape flags select the other configs.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Reference throughput published by the reference repo: none -> vs_baseline is null.
PUBLISHED_BASELINE = None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--method", default="ddp", choices=["single", "ddp", "fsdp", "mp"])
    ap.add_argument("--model_size", type=int, default=None)
    ap.add_argument("--ffn_dim", type=int, default=None)
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--batch_size", type=int, default=8)
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--comm", default="auto", choices=["auto", "fused", "native", "nccl"])
    ap.add_argument("--ops", default="native", choices=["native", "torch"])
    ap.add_argument("--recompute", action="store_true", help="reference-style activation recompute in backward")
    ap.add_argument("--mp_mode", default="ar", choices=["ar", "sp", "spf"],
                    help="MP form: ar = the reference's all-reduce form (default, the benchmarked one); sp / spf = sequence-parallel variants")
    ap.add_argument("--fsdp_gather", default="fused", choices=["dma", "fused"])
    ap.add_argument("--mp_input", default="sharded", choices=["full", "sharded"])
    ap.add_argument("--nvls", default="auto", choices=["auto", "off"])
    ap.add_argument("--also", default="auto",
                    help="extra strategies measured in the same process and reported under `also` (comma list of fsdp, mp, "
                         "ddp_recompute; auto = fsdp,mp when --method ddp; none)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--pool", type=int, default=3, help="distinct synthetic batches cycled through")
    return ap.parse_args()


def shape_defaults(args, method=None):
    method = method or args.method
    if method == "mp":          # BASELINE.json config 4: Llama-70B FFN shape, TP
        d, f, l = 8192, 28672, 16
    else:                       # configs 2/3: DDP / FSDP stack
        d, f, l = 4096, 16384, 32
    return args.model_size or d, args.ffn_dim or f, args.layers or l


def init_scale(d: int, f: int) -> float:
    """Variance-preserving init for the ReLU block: per-block gain^2 = D*F*s^4/2 = 1.  With the reference's fixed
    2e-2 (train_ffns.py:35) a 32-block d4096/f16384 stack gains 2.3x per block and overflows bf16 to NaN within a
    step - in BOTH arms - so the timed kernels would be chewing NaN/zero data."""
    return (2.0 / (d * f)) ** 0.25


# Upstream-gradient scale of the synthetic batches.  The reference's 0.1 (DLOSS_DX_COEF, train_ffns.py:30, "we may want
# to scale them depending on the model size & iterations") makes THIS stack diverge in ~4 steps in either arm: with
# no loss and no normalisation one SGD step moves every output by about lr*L*F*mean(a)^2*|sum_t dloss| ~ 7 sigma.  1e-4
# keeps activations and weights finite for hundreds of steps, so the timed kernels see ordinary N(0,1)-like data.
DLOSS_SCALE = 1e-4


def rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def make_host_pool(n, tokens, d, seed0):
    """Synthetic batches of the benchmark shape in PINNED host memory (bf16), like mock_data (:144-151)."""
    pool = []
    for i in range(n):
        g = torch.Generator().manual_seed(seed0 + i)
        x = torch.randn((tokens, d), generator=g, dtype=torch.float32).to(torch.bfloat16).pin_memory()
        dy = (DLOSS_SCALE * torch.randn((tokens, d), generator=g, dtype=torch.float32)).to(torch.bfloat16).pin_memory()
        pool.append((x, dy))
    return pool


def barrier_sync(world):
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_ranks(v: float, world, device) -> float:
    import torch.distributed as dist
    t = torch.tensor([v], dtype=torch.float32, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def emit(rec):
    print(json.dumps(rec), flush=True)


# ------------------------------------------------------------------------------------------------ shared
def base_config(method, world, d, f, l, args, global_batch):
    """The `config` block: IDENTICAL keys and values in both arms (what differs per arm goes to `impl_detail`)."""
    par = {"single": "single", "ddp": f"dp{world}", "fsdp": f"fsdp{world}", "mp": f"tp{world}"}[method]
    return {"model": f"ffn_stack_d{d}_f{f}_l{l}_relu", "global_batch": global_batch, "seq_len": args.seq_len,
            "batch_size_per_rank": args.batch_size, "parallelism": par,
            "l2": "per-step inputs (>= 256 MiB) and weights (>= 8 GiB) far exceed the 126 MB L2; batches cycle through a pool",
            "init": "variance-preserving scale (2/(D*F))^(1/4), same in both arms", "dloss_scale": DLOSS_SCALE}


def record(impl, method, world, args, value, ms, clocks, e2e, launches, config, detail):
    return {"metric": f"ffn_stack_{method}_tokens_per_sec", "value": value, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak" if method in ("ddp", "fsdp") else "strong", "vs_baseline": PUBLISHED_BASELINE,
            "dtype": "bf16", "data": "synthetic", "impl": impl, "config": config, "impl_detail": detail,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches}


def also_in_child(args, impl, method, idx, recompute=False):
    """Measure one more strategy in a FRESH process per rank (same launcher environment, its own rendezvous port), so that
    it starts from the same clean CUDA context / allocator state as a stand-alone `bench.py --method ...` run: measured in
    round 2, FSDP's end-to-end loop ran 13 % slower when it followed another engine in the same process.  Every rank spawns
    its child at the same point; rank 0 returns the child's JSON record."""
    import subprocess
    rank, world, _ = rank_world()
    # torchrun tells its workers to use the AGENT's rendezvous store (TORCHELASTIC_USE_AGENT_STORE): a child on another
    # port would wait for a store nobody serves.  Without those variables rank 0's child hosts its own store.
    env = {k: v for k, v in os.environ.items() if not k.startswith("TORCHELASTIC_")}
    if world > 1:
        env["MASTER_PORT"] = str(int(os.environ.get("MASTER_PORT", "29500")) + 101 + idx)
    env["B200_INIT_TIMEOUT_S"] = "180"
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", impl, "--method", method, "--also", "none", "--gpus", str(args.gpus),
           "--steps", str(args.steps), "--warmup", str(args.warmup), "--batch_size", str(args.batch_size), "--seq_len", str(args.seq_len),
           "--comm", args.comm, "--ops", args.ops, "--mp_mode", args.mp_mode, "--fsdp_gather", args.fsdp_gather,
           "--mp_input", args.mp_input, "--nvls", args.nvls, "--pool", str(args.pool)]
    if recompute:
        cmd.append("--recompute")
    if args.no_e2e:
        cmd.append("--no-e2e")
    try:
        p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        out, err, rc = p.stdout, p.stderr, p.returncode
    except subprocess.TimeoutExpired as e:
        out, err, rc = (e.stdout or ""), "timeout", -1
    if isinstance(out, bytes):
        out = out.decode(errors="replace")
    rec = None
    if rank == 0:
        for line in reversed(out.splitlines()):
            if line.startswith("{"):
                try:
                    rec = json.loads(line)
                    break
                except json.JSONDecodeError:
                    pass
        if rec is not None and "value" not in rec:
            rec = None
    # every rank must agree on whether the child run counts (a failed child falls back to an in-process measurement)
    ok = rc == 0 and (rank != 0 or rec is not None)
    if world > 1:
        import torch.distributed as dist
        flag = torch.tensor([1 if ok else 0], device=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
    if not ok:
        if rank == 0:
            print(f"bench: child run of {impl}/{method} failed (rc={rc}): {str(err)[-300:]} - measuring in-process instead", file=sys.stderr, flush=True)
        return "fallback"
    return rec


def also_methods(args):
    if args.also == "none":
        return []
    if args.also != "auto":
        return [m for m in args.also.split(",") if m]
    return ["fsdp", "mp"] if args.method == "ddp" else []


# ------------------------------------------------------------------------------------------------ ours
def ours_one(args, method, recompute):
    """One strategy of this framework: device-timed K steps, then the end-to-end loop.  Returns the record (rank 0's is
    the one printed).  Everything the run allocated is released before returning."""
    import gc
    from distributed_llm_code_samples_b200.config import TrainConfig
    from distributed_llm_code_samples_b200.engine import Engine, init_params_on_device
    from distributed_llm_code_samples_b200.ops import native as nv
    from distributed_llm_code_samples_b200.utils.clocks import ClockSampler

    rank, world, local = rank_world()
    d, f, l = shape_defaults(args, method)
    cfg = TrainConfig(num_steps=args.steps, batch_size=args.batch_size, seq_len=args.seq_len, layers=l, model_size=d,
                      ffn_dim=f, dtype="bf16", device="cuda", backend="nccl", impl=args.ops, comm=args.comm,
                      recompute=recompute, mp_mode=args.mp_mode, fsdp_gather=args.fsdp_gather, mp_input=args.mp_input, nvls=args.nvls)
    torch.cuda.set_device(local)
    params0 = init_params_on_device(cfg, torch.device("cuda", local), seed=1234, scale=init_scale(d, f))
    eng = Engine(cfg, method=method, layers_params=params0)
    del params0
    dev = eng.device
    tokens = cfg.tokens
    host_pool = make_host_pool(args.pool, tokens, d, 1000 * (rank if method in ("ddp", "fsdp") else 0))
    dev_pool = [(x.to(dev, non_blocking=True), dy.to(dev, non_blocking=True)) for x, dy in host_pool]
    torch.cuda.synchronize()

    # ---- device-timed: inputs staged on the device, K steps between CUDA events
    for i in range(args.warmup):
        eng.step(*dev_pool[i % len(dev_pool)])
    barrier_sync(world)
    sampler = ClockSampler(local).start() if rank == 0 else None
    launches0 = nv.launch_count
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(args.steps):
        eng.step(*dev_pool[i % len(dev_pool)])
    e.record()
    barrier_sync(world)
    launches = nv.launch_count - launches0
    ms = max_ranks(s.elapsed_time(e), world, dev)
    clocks = sampler.stop() if sampler is not None else None
    value = eng.tokens_per_step_global * args.steps / (ms * 1e-3)

    # ---- end to end: public API, pinned-host inputs every step, metric read back every step
    e2e = None
    if not args.no_e2e:
        def batches(n):
            for i in range(n):
                yield host_pool[i % len(host_pool)]
        for _ in eng.train(batches(max(1, args.warmup))):
            pass
        eng.h2d_bytes = eng.d2h_bytes = 0
        barrier_sync(world)
        t0 = time.perf_counter()
        metrics = list(eng.train(batches(args.steps)))
        barrier_sync(world)
        wall_ms = max_ranks((time.perf_counter() - t0) * 1e3, world, dev)
        e2e = {"value": eng.tokens_per_step_global * args.steps / (wall_ms * 1e-3), "unit": "tokens/s",
               "ms_per_step": wall_ms / args.steps, "h2d_bytes_per_step": eng.h2d_bytes // args.steps,
               "d2h_bytes_per_step": eng.d2h_bytes // args.steps, "last_metric": metrics[-1] if metrics else None}

    detail = {"activations": "recompute" if recompute else "kept",
              "gemms_per_block_step": 7 if recompute else 6,
              "comm": getattr(eng.trainer, "comm_name", "nccl" if world > 1 else "none"), "ops": args.ops,
              "torch_fallback_ops": getattr(eng.ops, "fallbacks", None),
              "peak_mem_gib": round(torch.cuda.max_memory_allocated(dev) / 2 ** 30, 2),     # torch's allocator only
              "device_mem_used_gib": round((lambda fr_to: (fr_to[1] - fr_to[0]) / 2 ** 30)(torch.cuda.mem_get_info(dev)), 2)}   # + symmetric blocks, NCCL, context
    rec = record("ours", method, world, args, value, ms, clocks, e2e, launches,
                 base_config(method if world > 1 else "single", world, d, f, l, args, eng.tokens_per_step_global), detail)
    eng.close()
    del eng, dev_pool, host_pool
    gc.collect()
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    return rec


def run_ours(args):
    rank, world, _ = rank_world()
    rec = ours_one(args, args.method, args.recompute)
    also = {}
    for m in also_methods(args):
        if world == 1 and m == "fsdp" and args.method == "ddp":
            # a world of one: FSDP and DDP are the same single-device loop on the same shape - not run twice
            also[m] = {"same_run_as": "ddp", "value": rec["value"], "unit": rec["unit"], "ms_per_step": rec["ms_per_step"],
                       "e2e": rec["e2e"], "clocks": rec["clocks"], "gpu_launches": rec["gpu_launches"], "config": rec["config"]}
            continue
        k = len(also)
        rm, rr = ("ddp", True) if m == "ddp_recompute" else (m, args.recompute)
        try:
            r = also_in_child(args, "ours", rm, k, rr)
            also[m] = ours_one(args, rm, rr) if r == "fallback" else r
        except Exception as ex:  # noqa: BLE001 - an extra strategy must never cost the headline line
            also[m] = {"error": f"{type(ex).__name__}: {ex}"[:400]}
    if also:
        rec["also"] = also
    if rank == 0:
        emit(rec)


# ------------------------------------------------------------------------------------------------ reference
def load_reference():
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "train_ffns.py")):
        return None, "baseline/_ref/train_ffns.py missing (reference not installed)"
    if not torch.cuda.is_available():
        return None, "no CUDA device"
    sys.path.insert(0, ref_dir)
    real_count = torch.cuda.device_count
    if real_count() == 1:       # the module refuses to import with exactly one visible GPU (train_ffns.py:26-27)
        torch.cuda.device_count = lambda: 2
    import train_ffns as ref    # UNMODIFIED reference module
    torch.cuda.device_count = real_count
    return ref, None


def reference_one(args, ref, method):
    """One strategy of the unmodified reference (its own per-rank function, its own H2D, its own collectives)."""
    import gc
    from distributed_llm_code_samples_b200.utils.clocks import ClockSampler
    rank, world, local = rank_world()
    d, f, l = shape_defaults(args, method)
    tokens = args.batch_size * args.seq_len
    dev = torch.device("cuda", local)

    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    full = []
    for _ in range(l):
        w1 = (init_scale(d, f) * torch.randn((f, d), generator=g, device=dev)).to(torch.bfloat16)
        w2 = (init_scale(d, f) * torch.randn((d, f), generator=g, device=dev)).to(torch.bfloat16)
        full.append([w1, w2])
    if method in ("ddp", "single") or world == 1:
        params = full
    elif method == "fsdp":
        params = [[p.chunk(world)[rank].contiguous() for p in lp] for lp in full]
    else:
        params = [[lp[0].chunk(world, dim=0)[rank].contiguous(), lp[1].chunk(world, dim=1)[rank].contiguous()] for lp in full]
    del full
    fn = {"single": ref.train_1gpu, "ddp": ref.train_process_ddp, "fsdp": ref.train_process_fsdp, "mp": ref.train_process_tp}[method]
    if world == 1:
        fn = ref.train_1gpu
    state = {"params": params}
    del params

    host_pool = make_host_pool(args.pool, tokens, d, 1000 * (rank if method in ("ddp", "fsdp") else 0))
    dev_pool = [(x.to(dev), dy.to(dev)) for x, dy in host_pool]

    def run_steps(n, pool):
        # the reference's loops pull batches from mock_data (:144); feed it the benchmark's bf16 batches
        def fake_mock_data(seeds, batch_size, model_size):
            for i in range(len(seeds)):
                yield pool[i % len(pool)]
        ref.mock_data = fake_mock_data
        seeds = torch.arange(n)
        if fn is ref.train_1gpu:
            state["params"] = fn(state["params"], seeds, tokens, d)      # out-of-place SGD returns new lists (:114)
        else:
            fn(local, state["params"], seeds, tokens, d)

    run_steps(args.warmup, dev_pool)
    barrier_sync(world)
    sampler = ClockSampler(local).start() if rank == 0 else None
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    run_steps(args.steps, dev_pool)
    e.record()
    barrier_sync(world)
    ms = max_ranks(s.elapsed_time(e), world, dev)
    clocks = sampler.stop() if sampler is not None else None
    per_step_tokens = tokens * (world if method in ("ddp", "fsdp") and world > 1 else 1)
    value = per_step_tokens * args.steps / (ms * 1e-3)

    e2e = None
    if not args.no_e2e:
        run_steps(1, host_pool)
        barrier_sync(world)
        t0 = time.perf_counter()
        run_steps(args.steps, host_pool)          # the reference's own .cuda(local_rank) does the H2D (:158)
        peek = state["params"][0][0][:5, :5].float().cpu()       # its final [:5,:5] read-back (:383-384)
        barrier_sync(world)
        wall_ms = max_ranks((time.perf_counter() - t0) * 1e3, world, dev)
        e2e = {"value": per_step_tokens * args.steps / (wall_ms * 1e-3), "unit": "tokens/s", "ms_per_step": wall_ms / args.steps,
               "h2d_bytes_per_step": 2 * tokens * d * 2, "d2h_bytes_per_step": int(peek.numel() * 4 / max(args.steps, 1))}

    detail = {"activations": "recompute", "gemms_per_block_step": 7,
              "comm": "NCCL via torch.distributed" if world > 1 else "none", "ops": "cuBLAS + ATen (torch)",
              "note": "unmodified baseline/_ref/train_ffns.py per-rank function; mock_data yields the bf16 benchmark batches",
              "final_param_finite": bool(torch.isfinite(state["params"][-1][0].float()).all().item()),
              "peak_mem_gib": round(torch.cuda.max_memory_allocated(dev) / 2 ** 30, 2),
              "device_mem_used_gib": round((lambda fr_to: (fr_to[1] - fr_to[0]) / 2 ** 30)(torch.cuda.mem_get_info(dev)), 2)}
    rec = record("reference", method, world, args, value, ms, clocks, e2e, 0,
                 base_config(method if world > 1 else "single", world, d, f, l, args, per_step_tokens), detail)
    del state, dev_pool, host_pool
    gc.collect()
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    return rec


def run_reference(args):
    ref, why = load_reference()
    if ref is None:
        emit({"impl": "reference", "unavailable": why})
        return
    import torch.distributed as dist
    rank, world, local = rank_world()
    ref.nGPUs = world           # it derives the world size from device_count at import (:25)
    torch.cuda.set_device(torch.device("cuda", local))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world)
    rec = reference_one(args, ref, args.method)
    also = {}
    for m in also_methods(args):
        if m == "ddp_recompute":
            continue                                  # the reference always recomputes (:63)
        if world == 1 and m == "fsdp" and args.method == "ddp":
            also[m] = {"same_run_as": "ddp", "value": rec["value"], "unit": rec["unit"], "ms_per_step": rec["ms_per_step"],
                       "e2e": rec["e2e"], "clocks": rec["clocks"], "gpu_launches": 0, "config": rec["config"]}
            continue
        try:
            r = also_in_child(args, "reference", m, len(also) + 10)
            also[m] = reference_one(args, ref, m) if r == "fallback" else r
        except Exception as ex:  # noqa: BLE001
            also[m] = {"error": f"{type(ex).__name__}: {ex}"[:400]}
    if also:
        rec["also"] = also
    if rank == 0:
        emit(rec)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if not torch.cuda.is_available():
        emit({"impl": args.impl, "unavailable": "no CUDA device on this box"})
        return 0
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
