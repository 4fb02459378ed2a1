#!/bin/bash
# Round 2, third call (N GPUs): the reworked fused kernels in the real multi-GPU tests, then bench.py with `also` records.
N=${1:-2}
O=gpurun_out/r2c3_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
step "multigpu tests" 600 python -m pytest tests/test_multigpu.py -q -m gpu > $O/mgpu.log 2>&1; tail -6 $O/mgpu.log | cut -c1-300
step "bench ours (ddp + also fsdp, mp)" 600 $TR --master-port 29531 bench.py --gpus $N --steps 4 --warmup 3 2> $O/bench_ours.err | grep "^{" > $O/bench_ours.json
step "bench reference (ddp + also fsdp, mp)" 600 $TR --master-port 29532 bench.py --impl reference --gpus $N --steps 4 --warmup 3 2> $O/bench_ref.err | grep "^{" > $O/bench_ref.json
python - <<PY
import json
for arm in ("ours", "ref"):
    try:
        d = json.load(open("$O/bench_%s.json" % arm))
        recs = [("ddp", d)] + list(d.get("also", {}).items())
        for m, r in recs:
            e = r.get("e2e") or {}
            print(arm, m, "n=$N", round(r["value"]), "tok/s", round(r["ms_per_step"], 2), "ms/step", (r.get("clocks") or {}).get("sm_mhz"), "MHz e2e", round(e.get("value", 0)), (r.get("impl_detail") or {}).get("comm", "")[:60])
    except Exception as ex:
        print(arm, "FAILED", ex); print(open("$O/bench_%s.err" % ("ours" if arm == "ours" else "ref")).read()[-1500:])
PY
