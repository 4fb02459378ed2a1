#!/bin/bash
# One ncu capture per kernel family (1 GPU).  Each case first runs plainly (exit 0 required), then once under ncu.
# Raw counter pages are exported on the box; only three .ncu-rep files travel back (size limit).
O=gpurun_out/r2ncu
mkdir -p $O
# (first, without any profiler: the new single-GPU tests and the isolated GEMM timings)
timeout 600 python -m pytest tests -q -m gpu --deselect tests/test_multigpu.py > $O/suite_1gpu.log 2>&1; echo "suite rc=$?"; tail -4 $O/suite_1gpu.log | cut -c1-300
timeout 200 python tools/gemm_diag.py 2>&1 | grep "^{\"case" | cut -c1-260 > $O/gemm_bench.jsonl; cat $O/gemm_bench.jsonl
KEEP="fused_dw G5m emu8_dw"
cap() {  # case, kernel regex, launches to skip
  local c=$1 k=$2 s=$3
  echo "== $c"
  python tools/ncu_gemm.py $c 5 > $O/plain_$c.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c 1 -f -o $O/prof_$c python tools/ncu_gemm.py $c 5 > $O/ncu_$c.log 2>&1
  echo "rc=$?"
  if [ -f $O/prof_$c.ncu-rep ]; then
    ncu -i $O/prof_$c.ncu-rep --page raw --csv > $O/prof_$c.raw.csv 2>/dev/null
    case " $KEEP " in *" $c "*) ;; *) rm -f $O/prof_$c.ncu-rep ;; esac
  else
    tail -5 $O/ncu_$c.log; tail -3 $O/plain_$c.log
  fi
}
cap G1m gemm_tcgen05 4
cap G5m gemm_tcgen05 4
cap G7 gemm_tcgen05 3
cap G6 gemm_tcgen05 3
cap G2 gemm_tcgen05 3
cap fused_dw gemm_tcgen05 3
cap fused_adam gemm_tcgen05 3
cap fused_mp_1cta gemm_tcgen05 3
cap emu8_dw gemm_tcgen05_emu 3
cap emu8_mp gemm_tcgen05_emu 3
cap gather gemm_tcgen05 3
cap sgd sgd_kernel 3
cap adam adam_kernel 3
cap sumsq sumsq_kernel 3
cap philox randn_kernel 3
cap colsum colsum_partial 3
cap coll_ar coll_kernel 3
cap f32 gemm_tcgen05 3
ls -la $O | head -60
