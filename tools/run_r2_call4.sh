#!/bin/bash
# Round 2, fourth call (N GPUs): north-star block on the fused trainers (bias + GELU + residual, SGD / Adam), the new
# elementwise kernels, and the end-to-end-loop probe.
N=${1:-2}
O=gpurun_out/r2c4_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
step "1-GPU: new kernels + north-star block" 300 python -m pytest tests/test_gemm_gpu.py tests/test_train_gpu.py -q -m gpu -k "adam or fp32_param or north_star or cuda_graph or prefetcher or elementwise" > $O/new_1gpu.log 2>&1; tail -5 $O/new_1gpu.log | cut -c1-300
step "multi-GPU: north star + native-collective trainers" 600 python -m pytest tests/test_multigpu.py -q -m gpu -k "north_star or unfused_trainers" > $O/mgpu_north_star.log 2>&1; tail -6 $O/mgpu_north_star.log | cut -c1-400
step "e2e probe fsdp" 200 $TR --master-port 29541 tools/e2e_probe.py --method fsdp 2> $O/e2e_fsdp.err | grep "^{" | tee $O/e2e_fsdp.json
step "e2e probe ddp" 200 $TR --master-port 29542 tools/e2e_probe.py --method ddp 2> $O/e2e_ddp.err | grep "^{" | tee $O/e2e_ddp.json
step "e2e probe mp full" 200 $TR --master-port 29543 tools/e2e_probe.py --method mp 2> $O/e2e_mp.err | grep "^{" | tee $O/e2e_mp.json
step "e2e probe mp sharded" 200 $TR --master-port 29544 tools/e2e_probe.py --method mp --mp_input sharded 2> $O/e2e_mp_sh.err | grep "^{" | tee $O/e2e_mp_sh.json
for f in $O/*.err; do [ -s $f ] && grep -iE "error|Traceback" -A12 $f | tail -30; done
