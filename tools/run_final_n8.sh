#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== pytest multigpu"; timeout 500 python -m pytest tests/test_multigpu.py -q -m gpu > gpurun_out/pytest_mgpu_n$N.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/pytest_mgpu_n$N.log
for v in "2 4" "0 4" "1 4" "1 2"; do set -- $v
  echo "== mp fused: overlap=$1 side=$2"; FUSED_BENCH_ONLY=mp FUSED_BENCH_TAG=_ov$1_sw$2 B200_FUSED_OVERLAP=$1 B200_FUSED_SIDE_WARPS=$2 timeout 200 $TR --master-port 29514 tools/fused_bench.py 2>&1 | grep "^{" | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['fused_gemm_allreduce_ms'], d['gemm_nccl_allreduce_ms'], d['local_gemm_ms'], d['roofline_frac_fused'])"
done
for m in ddp fsdp mp; do
  timeout 400 $TR --master-port 29511 bench.py --method $m --gpus $N --steps 4 --warmup 3 2> gpurun_out/bench_ours_${m}_n$N.err | grep "^{" > gpurun_out/bench_ours_${m}_n$N.json
  python -c "
import json
try:
    d=json.load(open('gpurun_out/bench_ours_${m}_n$N.json')); print('ours $m', round(d['value']), round(d['ms_per_step'],1), d['clocks']['sm_mhz'], round(d['e2e']['value']), d['e2e']['last_metric'])
except Exception as e:
    print('FAILED', e); print(open('gpurun_out/bench_ours_${m}_n$N.err').read()[-800:])
"
done
