#!/bin/bash
# 2 GPUs: the reference's own method-0 run (1 GPU, DDP, FSDP, TP from the same initial weights, DDP == FSDP check) on the fused
# bf16 paths and in fp32 (tcgen05 split-operand GEMMs + NCCL), with the difference reported in ulps.
O=gpurun_out/r2c9
mkdir -p $O
for dt in bf16 fp32; do
  echo "== train_ffns -m 0 --dtype $dt"
  timeout -k 15 240 python train_ffns.py -s 8 -bs 4 -n 1024 -l 2 -d 2048 -m 0 -r 7 --dtype $dt --lr 1e-3 > $O/train_m0_$dt.log 2>&1; echo "rc=$?"
  grep -E "takes|SoftAssertion|DDP vs FSDP|tokens/s" $O/train_m0_$dt.log | cut -c1-220 | head -12
done
