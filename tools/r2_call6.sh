#!/bin/bash
mkdir -p gpurun_out/c6
for sk in 7 0; do
  echo "== CF_TC2_SKIP=$sk" >> gpurun_out/c6/sweep.log
  CF_TC2_SKIP=$sk timeout 300 python tools/build_sweep.py >> gpurun_out/c6/sweep.log 2>&1
done
echo "== v1 (W%8 forced off via CF_FORCE_V1)" >> gpurun_out/c6/sweep.log
CF_FORCE_V1=1 timeout 300 python tools/build_sweep.py >> gpurun_out/c6/sweep.log 2>&1
grep -E "==|build_sweep" gpurun_out/c6/sweep.log
