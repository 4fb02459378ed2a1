#!/bin/bash
# Round 2, last call (1 GPU): the whole single-GPU suite on the final code, then ONE compute-sanitizer tool (memcheck) on the
# smallest tcgen05 / TMA case.
O=gpurun_out/r2c12
mkdir -p $O
timeout -k 10 150 python -m pytest tests -q -m gpu --deselect tests/test_multigpu.py > $O/suite_1gpu.log 2>&1; echo "suite rc=$?"; tail -3 $O/suite_1gpu.log | cut -c1-300
timeout -k 10 120 compute-sanitizer --tool memcheck --log-file $O/sanitizer_memcheck.log python -m pytest tests/test_gemm_gpu.py -x -q -m gpu -k "test_gemm_tn and 128-256-64" > $O/sanitizer_memcheck.out 2>&1; echo "memcheck rc=$?"; tail -4 $O/sanitizer_memcheck.log; tail -2 $O/sanitizer_memcheck.out
