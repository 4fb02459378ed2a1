#!/bin/bash
# One compute-sanitizer tool per call, on the smallest case that exercises the TMA/tcgen05 pipeline and the fused
# epilogue in a world of one.  usage: tools/sanitize.sh memcheck|racecheck|synccheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 600 compute-sanitizer --tool "$TOOL" --log-file gpurun_out/sanitizer_$TOOL.log \
  python -m pytest tests/test_gemm_gpu.py -x -q -m gpu -k "test_gemm_tn and 128-256-64" > gpurun_out/sanitizer_$TOOL.out 2>&1
echo "rc=$?"; tail -5 gpurun_out/sanitizer_$TOOL.log
