#!/bin/bash
# Round 2, first multi-GPU call: run what round 1 wrote but never executed, cheapest first, one test per process so a
# hang or trap is attributed.  Logs under gpurun_out/r2c1/ (copied to profiles/tests/ once read).
#   gpurun --gpus 2 --timeout 1150 -- 'bash tools/run_r2_call1.sh 2'
N=${1:-2}
O=gpurun_out/r2c1
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
export B200_QUEUED_TESTS=1
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }

nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.csv 2>&1
nvidia-smi topo -m > $O/topo.txt 2>&1

# 1. the validated (non-queued) multi-GPU tests: a retained PASSING log of the fused-vs-NCCL-vs-oracle comparison
B200_QUEUED_TESTS=0 step "multigpu standard" 420 python -m pytest tests/test_multigpu.py -q -m gpu -x > $O/mgpu_standard.log 2>&1; tail -4 $O/mgpu_standard.log

# 2. is NVLS there at all?
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL step "nvls probe" 150 $TR --master-port 29526 tools/nvls_probe.py > $O/nvls_probe.log 2>&1; grep -iE "nvls|multicast|probe:" $O/nvls_probe.log | head -8

# 3. queued tests, one per process
for k in barrier_and_flag_protocol_stress vmm_multicast_all_reduce "fused_trainer_matches_nccl_trainer and mp_spf" fused_adam_matches_unfused_adam fused_mp_with_gelu_and_residual fsdp_with_in_kernel_forward_gather unfused_trainers_on_native_collectives; do
  tag=$(echo "$k" | tr ' ' '_')
  step "multigpu: $k" 300 python -m pytest tests/test_multigpu.py -q -m gpu -k "$k" > $O/mgpu_$tag.log 2>&1; tail -4 $O/mgpu_$tag.log | cut -c1-300
done

# 4. the fused dW kernel in the DDP step at N ranks: polling / side-warp variants (the 1 -> 2 drop is the biggest)
for v in "0 4" "1 4" "0 2"; do
  set -- $v
  B200_FUSED_POLL=$1 B200_FUSED_SIDE_WARPS=$2 step "bench ddp poll=$1 side=$2" 240 $TR --master-port 29522 bench.py --method ddp --gpus $N --steps 4 --warmup 3 --no-e2e 2> $O/bench_ddp_p$1_s$2.err | grep "^{" > $O/bench_ddp_p$1_s$2.json
  python - <<PY
import json
try:
    d = json.load(open("$O/bench_ddp_p$1_s$2.json")); print("ddp poll=$1 side=$2", round(d["value"]), "tok/s", round(d["ms_per_step"], 2), "ms/step", d["clocks"]["sm_mhz"], "MHz")
except Exception as e:
    print("ddp poll=$1 side=$2 FAILED", e); print(open("$O/bench_ddp_p$1_s$2.err").read()[-800:])
PY
done
