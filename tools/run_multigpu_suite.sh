#!/bin/bash
# Full multi-GPU evidence run: tools/run_multigpu_suite.sh N   (inside gpurun --gpus N)
N=${1:-2}
STEPS=${2:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== pytest multigpu"; timeout 600 python -m pytest tests/test_multigpu.py -q -m gpu > gpurun_out/pytest_mgpu_n$N.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_mgpu_n$N.log
echo "== fused per-kernel bench"; timeout 400 $TR --master-port 29514 tools/fused_bench.py 2>&1 | grep "^{" | cut -c1-600
echo "== collective sweep"; timeout 400 $TR --master-port 29515 tools/coll_sweep.py 2>&1 | grep "^{" | tail -4 | cut -c1-400
for m in ddp fsdp mp; do
  for impl in reference ours; do
    timeout 600 $TR --master-port 29511 bench.py --impl $impl --method $m --gpus $N --steps $STEPS --warmup 3 2> gpurun_out/bench_${impl}_${m}_n$N.err | grep "^{" > gpurun_out/bench_${impl}_${m}_n$N.json
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_${impl}_${m}_n$N.json"))
    print("$impl $m n=$N", round(d["value"]), "tok/s", round(d["ms_per_step"], 1), "ms/step clk", d["clocks"]["sm_mhz"], "e2e", round(d["e2e"]["value"]) if d.get("e2e") else None, d["config"].get("comm", "")[:50])
except Exception as e:
    print("$impl $m n=$N FAILED", e)
    print(open("gpurun_out/bench_${impl}_${m}_n$N.err").read()[-1500:])
PY
  done
done
