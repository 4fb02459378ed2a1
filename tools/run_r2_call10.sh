#!/bin/bash
# Round 2, closing 2-GPU call: new single-GPU tests, bench.py with the final defaults (FSDP gather inside the GEMMs, MP sharded
# input staging, NVLS), and the reference's method-0 run (DDP == FSDP check) in bf16 and fp32.
N=${1:-2}
O=gpurun_out/r2c10_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
step "1-GPU: transpose, fp32 mode, masks, smoke" 200 python -m pytest tests/test_gemm_gpu.py tests/test_train_gpu.py -q -m gpu -k "transpose or fp32 or f32 or bitmask or smoke or north_star" > $O/new_1gpu.log 2>&1; tail -4 $O/new_1gpu.log | cut -c1-300
step "multi-GPU: fsdp (dma + in-kernel gather), mp" 300 python -m pytest tests/test_multigpu.py -q -m gpu -k "(fused_trainer_matches and (fsdp or (mp and not mp_sp))) or fsdp_with_in_kernel" > $O/mgpu.log 2>&1; tail -4 $O/mgpu.log | cut -c1-300
step "bench ours, final defaults" 300 $TR --master-port 29561 bench.py --gpus $N --steps 4 --warmup 3 2> $O/bench_ours.err | grep "^{" > $O/bench_ours.json
python - <<PY
import json
try:
    d = json.load(open("$O/bench_ours.json"))
    for m, r in [("ddp", d)] + list(d.get("also", {}).items()):
        e = r.get("e2e") or {}; det = r.get("impl_detail") or {}
        print("ours", m, "n=$N", round(r.get("value", 0)), "tok/s", round(r.get("ms_per_step", 0), 2), "ms/step e2e", round(e.get("value", 0)), round(e.get("ms_per_step", 0), 2), "h2d", e.get("h2d_bytes_per_step"), "mem", det.get("peak_mem_gib"), det.get("device_mem_used_gib"), "fb", det.get("torch_fallback_ops"), det.get("comm", "")[:48], r.get("error", ""))
except Exception as ex:
    print("ours FAILED", ex); print(open("$O/bench_ours.err").read()[-1200:])
PY
grep -h "bench: child" $O/*.err | head -3
bash tools/run_r2_call9.sh
