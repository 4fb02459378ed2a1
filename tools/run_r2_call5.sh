#!/bin/bash
# Round 2, fifth call (N GPUs): NVLS - in-switch reduction inside the fused MP all-reduce and in the standalone collectives -
# plus bench.py with its `also` strategies measured in fresh child processes.
N=${1:-2}
O=gpurun_out/r2c5_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
step "multi-GPU: collectives (NVLS staging), vmm, stress, fused MP (NVLS)" 600 python -m pytest tests/test_multigpu.py -q -m gpu -k "native_collectives or vmm or stress or (fused_trainer_matches and mp) or (north_star and mp and sgd) or gelu_and_residual" > $O/mgpu_nvls.log 2>&1; tail -6 $O/mgpu_nvls.log | cut -c1-400
FUSED_BENCH_ONLY=mp FUSED_BENCH_TAG=_nvls step "fused_bench mp (push vs NVLS)" 240 $TR --master-port 29551 tools/fused_bench.py 2> $O/fused_bench.err | grep "^{" | cut -c1-800
step "coll sweep (NVLS vs P2P vs NCCL)" 300 $TR --master-port 29552 tools/coll_sweep.py 268435456 2> $O/coll_sweep.err | grep "^{" | cut -c1-420
for nv in auto off; do
  step "bench mp nvls=$nv" 300 $TR --master-port 29553 bench.py --method mp --nvls $nv --gpus $N --steps 4 --warmup 3 --no-e2e 2> $O/bench_mp_$nv.err | grep "^{" > $O/bench_mp_$nv.json
  python - <<PY
import json
try:
    d = json.load(open("$O/bench_mp_$nv.json")); print("mp nvls=$nv", round(d["value"]), "tok/s", round(d["ms_per_step"], 2), "ms/step", d["clocks"]["sm_mhz"], "MHz", d["impl_detail"]["comm"][:70])
except Exception as e:
    print("mp nvls=$nv FAILED", e); print(open("$O/bench_mp_$nv.err").read()[-1500:])
PY
done
step "bench ours (ddp + children fsdp, mp)" 900 $TR --master-port 29554 bench.py --gpus $N --steps 4 --warmup 3 2> $O/bench_ours.err | grep "^{" > $O/bench_ours.json
python - <<PY
import json
try:
    d = json.load(open("$O/bench_ours.json"))
    for m, r in [("ddp", d)] + list(d.get("also", {}).items()):
        e = r.get("e2e") or {}
        print("ours", m, "n=$N", round(r.get("value", 0)), "tok/s", round(r.get("ms_per_step", 0), 2), "ms/step e2e", round(e.get("value", 0)), r.get("error", ""), r.get("stderr_tail", "")[-300:])
except Exception as ex:
    print("ours FAILED", ex); print(open("$O/bench_ours.err").read()[-1500:])
PY
cp gpurun_out/fused_bench_n${N}_nvls.jsonl gpurun_out/coll_sweep_n$N.jsonl $O/ 2>/dev/null
for f in $O/*.err; do [ -s $f ] && grep -iE "Error|Traceback" -A8 $f | tail -24; done
