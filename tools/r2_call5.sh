#!/bin/bash
mkdir -p gpurun_out/c5
for sk in 7 15 8; do
  echo "== CF_TC2_SKIP=$sk" >> gpurun_out/c5/skip.log
  CF_TC2_SKIP=$sk timeout 200 python tools/kbench.py cfg2 --graph --only build_bf16,build_bf16x3 >> gpurun_out/c5/skip.log 2>&1
  CF_TC2_SKIP=$sk timeout 200 python tools/kbench.py cfg4 --graph --only build_bf16 >> gpurun_out/c5/skip.log 2>&1
done
grep -E "==|corr_build" gpurun_out/c5/skip.log
