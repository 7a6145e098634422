#!/bin/bash
mkdir -p gpurun_out/c4
for sk in 0 7; do
CF_TC2_SKIP=$sk timeout 200 python tools/kbench.py cfg2 --only build_bf16 --reps 5 > gpurun_out/c4/plain_$sk.log 2>&1 &&
CF_TC2_SKIP=$sk timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c4/launches_$sk.csv python tools/kbench.py cfg2 --only build_bf16 --reps 5 > gpurun_out/c4/ncu_$sk.log 2>&1
done
tail -12 gpurun_out/c4/launches_0.csv | cut -d, -f5,15 
tail -12 gpurun_out/c4/launches_7.csv | cut -d, -f5,15
