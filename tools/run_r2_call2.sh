#!/bin/bash
# Round 2, second call (1 GPU): the whole single-GPU suite (incl. the multi-rank EMULATION tests and the benchmark-size
# numerics), then A/B of the fused dW kernel variants and of the mbarrier suspend-hint build.
O=gpurun_out/r2c2
mkdir -p $O
export B200_QUEUED_TESTS=1
exec 3>&1
step() { echo "== $1" >&3; local t0=$SECONDS; shift; timeout -k 15 "$@"; echo "rc=$? ($((SECONDS-t0)) s)" >&3; }
V=distributed_llm_code_samples_b200/_native/variants
step "emu tests" 600 python -m pytest tests/test_emu_gpu.py -q -m gpu -x > $O/emu.log 2>&1; tail -5 $O/emu.log | cut -c1-300
step "1-GPU suite" 900 python -m pytest tests -q -m gpu --deselect tests/test_emu_gpu.py > $O/suite_1gpu.log 2>&1; tail -8 $O/suite_1gpu.log | cut -c1-300
step "fused_ab base" 200 python tools/fused_ab.py base 2>&1 | grep "^{" | cut -c1-600
for t in hint20k hint1k; do
  B200_NATIVE_LIB=$V/$t/libb200ffn_$t.so step "fused_ab $t" 200 python tools/fused_ab.py $t 2>&1 | grep "^{" | cut -c1-600
done
step "gemm bench base" 200 python tools/gemm_diag.py 2>&1 | grep "^{" | cut -c1-300
GEMM_BENCH_TAG=_hint20k B200_NATIVE_LIB=$V/hint20k/libb200ffn_hint20k.so step "gemm bench hint20k" 200 python tools/gemm_diag.py 2>&1 | grep "^{" | cut -c1-300
for t in base hint20k; do
  L=""; [ $t != base ] && L=$V/$t/libb200ffn_$t.so
  B200_NATIVE_LIB=$L step "bench n1 $t" 300 python bench.py --steps 6 --warmup 3 --no-e2e 2> $O/bench_n1_$t.err | grep "^{" > $O/bench_n1_$t.json
  python - <<PY
import json
try:
    d = json.load(open("$O/bench_n1_$t.json")); print("n1 $t", round(d["value"]), "tok/s", round(d["ms_per_step"], 2), "ms/step", d["clocks"]["sm_mhz"], "MHz")
except Exception as e:
    print("n1 $t FAILED", e); print(open("$O/bench_n1_$t.err").read()[-800:])
PY
done
cp gpurun_out/fused_ab_*.jsonl gpurun_out/gemm_bench_ffn*.jsonl $O/ 2>/dev/null
