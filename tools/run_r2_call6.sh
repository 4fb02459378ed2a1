#!/bin/bash
# Round 2, sixth call (N GPUs): bench.py both arms with the `also` strategies in child processes (strict timeouts), and the
# fused Adam-state resume test.
N=${1:-2}
O=gpurun_out/r2c6_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
step "bench ours (ddp + children fsdp, mp)" 420 $TR --master-port 29561 bench.py --gpus $N --steps 4 --warmup 3 2> $O/bench_ours.err | grep "^{" > $O/bench_ours.json
step "bench reference (ddp + children fsdp, mp)" 420 $TR --master-port 29571 bench.py --impl reference --gpus $N --steps 4 --warmup 3 2> $O/bench_ref.err | grep "^{" > $O/bench_ref.json
python - <<PY
import json
for arm in ("ours", "ref"):
    try:
        d = json.load(open("$O/bench_%s.json" % arm))
        for m, r in [("ddp", d)] + list(d.get("also", {}).items()):
            e = r.get("e2e") or {}
            print(arm, m, "n=$N", round(r.get("value", 0)), "tok/s", round(r.get("ms_per_step", 0), 2), "ms/step e2e", round(e.get("value", 0)), round(e.get("ms_per_step", 0), 2), (r.get("impl_detail") or {}).get("comm", "")[:50])
    except Exception as ex:
        print(arm, "FAILED", ex); print(open("$O/bench_%s.err" % arm).read()[-1200:])
PY
grep -h "bench: child" $O/*.err | head
step "multi-GPU: fused Adam state resume + native-collective trainers" 400 python -m pytest tests/test_multigpu.py -q -m gpu -k "state_resumes or unfused_trainers" > $O/mgpu_resume.log 2>&1; tail -5 $O/mgpu_resume.log | cut -c1-300
