#!/bin/bash
# First GPU call of the next round: run everything that was written while no GPU was available, cheapest and safest first.
#   gpurun --gpus 2 --timeout 1500 -- 'bash tools/run_queued_validation.sh 2'
# Every step has its own timeout (a hang must not take the box down) and logs under gpurun_out/queued/.
N=${1:-2}
O=gpurun_out/queued
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
export B200_QUEUED_TESTS=1
exec 3>&1                                  # step() reports on the real stdout even when the command's output is redirected
step() { echo "== $1" >&3; shift; timeout -k 15 "$@"; echo "rc=$?" >&3; }   # timeout signals the whole process group, ranks included

# 1. single GPU: CUDA-graph replay (bitwise vs eager), then the whole 1-GPU suite with the queued tests on
step "graph replay test" 300 python -m pytest tests/test_train_gpu.py -q -m gpu -k cuda_graph > $O/graph.log 2>&1; tail -3 $O/graph.log
step "1-GPU suite (queued on)" 900 python -m pytest tests -q -m gpu --deselect tests/test_multigpu.py > $O/suite_1gpu.log 2>&1; tail -3 $O/suite_1gpu.log

# 2a. is NVLS there at all?  (decides whether the vmm test below can do more than the peer-table half)
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL step "nvls probe" 200 $TR --master-port 29526 tools/nvls_probe.py > $O/nvls_probe.log 2>&1; grep -iE "nvls|multicast|probe:" $O/nvls_probe.log | head -12

# 2. multi GPU, one queued test per process so that a hang / trap is attributed
for k in barrier_and_flag_protocol_stress unfused_trainers_on_native_collectives fused_adam_matches_unfused_adam fused_mp_with_gelu_and_residual vmm_multicast_all_reduce "fused_trainer_matches_nccl_trainer and mp_spf" fsdp_with_in_kernel_forward_gather; do
  tag=$(echo "$k" | tr ' ' '_')
  step "multigpu: $k" 400 python -m pytest tests/test_multigpu.py -q -m gpu -k "$k" > $O/mgpu_$tag.log 2>&1; tail -4 $O/mgpu_$tag.log
done

# 3. per-kernel: all-gather + GEMM fused vs its unfused forms (only after the mp_spf test above passed)
FUSED_BENCH_ONLY=ag FUSED_BENCH_TAG=_ag step "fused_bench ag" 300 $TR --master-port 29523 tools/fused_bench.py 2>&1 | grep "^{" | cut -c1-700

# 4. MP forms, same config: ar (validated) vs sp (validated) vs spf (new); then the poll back-off on DDP
for mode in ar sp spf; do
  step "bench mp $mode" 500 $TR --master-port 29521 bench.py --method mp --mp_mode $mode --gpus $N --steps 4 --warmup 3 --no-e2e 2> $O/bench_mp_$mode.err | grep "^{" > $O/bench_mp_$mode.json
  python -c "
import json
try:
    d=json.load(open('$O/bench_mp_$mode.json')); print('mp $mode', round(d['value']), 'tok/s', round(d['ms_per_step'],2), 'ms/step', d['clocks']['sm_mhz'], 'MHz')
except Exception as e:
    print('mp $mode FAILED', e); print(open('$O/bench_mp_$mode.err').read()[-1200:])
"
done
for poll in 0 1; do
  B200_FUSED_POLL=$poll step "bench ddp poll=$poll" 500 $TR --master-port 29522 bench.py --method ddp --gpus $N --steps 4 --warmup 3 --no-e2e 2> $O/bench_ddp_poll$poll.err | grep "^{" > $O/bench_ddp_poll$poll.json
  python -c "
import json
try:
    d=json.load(open('$O/bench_ddp_poll$poll.json')); print('ddp poll=$poll', round(d['value']), 'tok/s', round(d['ms_per_step'],2), 'ms/step', d['clocks']['sm_mhz'], 'MHz')
except Exception as e:
    print('ddp poll=$poll FAILED', e); print(open('$O/bench_ddp_poll$poll.err').read()[-1200:])
"
done
# 6. MP end to end: full-batch staging on every rank vs 1/world staged + NVLink all-gather
for sh in 0 1; do
  B200_MP_SHARDED_INPUT=$sh step "bench mp e2e sharded_input=$sh" 500 $TR --master-port 29524 bench.py --method mp --gpus $N --steps 4 --warmup 3 2> $O/bench_mp_e2e_sh$sh.err | grep "^{" > $O/bench_mp_e2e_sh$sh.json
  python -c "
import json
try:
    d=json.load(open('$O/bench_mp_e2e_sh$sh.json')); print('mp sharded_input=$sh device', round(d['value']), 'e2e', round(d['e2e']['value']), 'h2d/step', d['e2e']['h2d_bytes_per_step'])
except Exception as e:
    print('mp sharded_input=$sh FAILED', e); print(open('$O/bench_mp_e2e_sh$sh.err').read()[-1200:])
"
done
# 7. FSDP: copy-engine prefetch (validated) vs forward all-gather inside the GEMMs
for fg in 0 1; do
  B200_FSDP_FUSED_GATHER=$fg step "bench fsdp fused_gather=$fg" 500 $TR --master-port 29525 bench.py --method fsdp --gpus $N --steps 4 --warmup 3 --no-e2e 2> $O/bench_fsdp_fg$fg.err | grep "^{" > $O/bench_fsdp_fg$fg.json
  python -c "
import json
try:
    d=json.load(open('$O/bench_fsdp_fg$fg.json')); print('fsdp fused_gather=$fg', round(d['value']), 'tok/s', round(d['ms_per_step'],2), 'ms/step')
except Exception as e:
    print('fsdp fused_gather=$fg FAILED', e); print(open('$O/bench_fsdp_fg$fg.err').read()[-1200:])
"
done
