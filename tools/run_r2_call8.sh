#!/bin/bash
# Round 2, the one 8-GPU call: numerics of the fused paths at 8 ranks (incl. NVLS), the benchmark with its child strategies,
# A/B of the optional paths, per-kernel and collective sweeps.  Every step has a strict timeout.
N=${1:-8}
O=gpurun_out/r2c8_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
one() {  # tag, extra bench flags...: one strategy, no children
  local tag=$1; shift
  step "bench $tag" 150 $TR --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 4 --warmup 3 --also none "$@" 2> $O/bench_$tag.err | grep "^{" > $O/bench_$tag.json
  python - <<PY
import json
try:
    d = json.load(open("$O/bench_$tag.json")); e = d.get("e2e") or {}
    print("$tag", round(d["value"]), "tok/s", round(d["ms_per_step"], 2), "ms/step", (d.get("clocks") or {}).get("sm_mhz"), "MHz e2e", round(e.get("value", 0)), round(e.get("ms_per_step", 0), 2), "h2d", e.get("h2d_bytes_per_step"), d["impl_detail"]["comm"][:60])
except Exception as ex:
    print("$tag FAILED", ex); print(open("$O/bench_$tag.err").read()[-800:])
PY
}
nvidia-smi topo -m > $O/topo.txt 2>&1
step "multi-GPU tests at $N ranks" 360 python -m pytest tests/test_multigpu.py -q -m gpu -k "(fused_trainer_matches and (ddp or mp_spf or (mp and not mp_sp))) or fsdp_with_in_kernel or gradient_sum" > $O/mgpu.log 2>&1; tail -5 $O/mgpu.log | cut -c1-400
step "bench ours (ddp + children fsdp, mp)" 330 $TR --master-port 29581 bench.py --gpus $N --steps 5 --warmup 3 2> $O/bench_ours.err | grep "^{" > $O/bench_ours.json
python - <<PY
import json
try:
    d = json.load(open("$O/bench_ours.json"))
    for m, r in [("ddp", d)] + list(d.get("also", {}).items()):
        e = r.get("e2e") or {}
        print("ours", m, "n=$N", round(r.get("value", 0)), "tok/s", round(r.get("ms_per_step", 0), 2), "ms/step", (r.get("clocks") or {}).get("sm_mhz"), "MHz e2e", round(e.get("value", 0)), round(e.get("ms_per_step", 0), 2), (r.get("impl_detail") or {}).get("comm", "")[:60])
except Exception as ex:
    print("ours FAILED", ex); print(open("$O/bench_ours.err").read()[-1200:])
PY
one mp_sharded_input --method mp --mp_input sharded
one mp_nvls_off --method mp --nvls off --no-e2e
one mp_spf --method mp --mp_mode spf --no-e2e
one fsdp_fused_gather --method fsdp --fsdp_gather fused --no-e2e
one ddp_recompute --method ddp --recompute --no-e2e
one fsdp_recompute --method fsdp --recompute --no-e2e
FUSED_BENCH_ONLY=mp FUSED_BENCH_TAG=_r2 step "fused_bench mp (push vs NVLS)" 120 $TR --master-port 29591 tools/fused_bench.py 2> $O/fused_bench.err | grep "^{" | cut -c1-900
step "coll sweep" 200 $TR --master-port 29592 tools/coll_sweep.py 2> $O/coll_sweep.err | grep "^{" | cut -c1-420
cp gpurun_out/fused_bench_n${N}_r2.jsonl gpurun_out/coll_sweep_n$N.jsonl $O/ 2>/dev/null
for f in $O/*.err; do [ -s $f ] && grep -iE "Error|Traceback" -A6 $f | tail -14; done
