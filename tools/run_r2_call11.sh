#!/bin/bash
# Round 2, 4-GPU check (never run before): fused DDP / MP / FSDP in-kernel gather numerics, then bench.py with children.
N=${1:-4}
O=gpurun_out/r2c11_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
step "multi-GPU tests at $N ranks" 200 python -m pytest tests/test_multigpu.py -q -m gpu -k "(fused_trainer_matches and (ddp or (mp and not mp_sp))) or fsdp_with_in_kernel or gradient_sum" > $O/mgpu.log 2>&1; tail -4 $O/mgpu.log | cut -c1-300
step "bench ours" 240 $TR --master-port 29571 bench.py --gpus $N --steps 4 --warmup 3 2> $O/bench_ours.err | grep "^{" > $O/bench_ours.json
python - <<PY
import json
try:
    d = json.load(open("$O/bench_ours.json"))
    for m, r in [("ddp", d)] + list(d.get("also", {}).items()):
        e = r.get("e2e") or {}; det = r.get("impl_detail") or {}
        print("ours", m, "n=$N", round(r.get("value", 0)), "tok/s", round(r.get("ms_per_step", 0), 2), "ms/step e2e", round(e.get("value", 0)), round(e.get("ms_per_step", 0), 2), det.get("comm", "")[:48], r.get("error", ""))
except Exception as ex:
    print("ours FAILED", ex); print(open("$O/bench_ours.err").read()[-1200:])
PY
