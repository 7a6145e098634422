#!/bin/bash
set -x
mkdir -p gpurun_out/c2
for sk in 0 1 2 4 6 7; do
  echo "== CF_TC2_SKIP=$sk" >> gpurun_out/c2/skip.log
  CF_TC2_SKIP=$sk timeout 200 python tools/kbench.py cfg2 --graph --only build_bf16 >> gpurun_out/c2/skip.log 2>&1
done
cat gpurun_out/c2/skip.log
timeout 300 python tools/kbench.py cfg2 --only build_bf16 --reps 6 > gpurun_out/c2/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:corr_tc2 -s 2 -c 1 -o gpurun_out/c2/tc2 python tools/kbench.py cfg2 --only build_bf16 --reps 6 > gpurun_out/c2/ncu.log 2>&1
tail -5 gpurun_out/c2/ncu.log
