#!/bin/bash
mkdir -p gpurun_out/c8
timeout 300 python tools/kbench.py cfg2 --only build_bf16 --reps 6 > gpurun_out/c8/plain1.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:corr_tc2 -s 2 -c 1 -o gpurun_out/c8/tc2 python tools/kbench.py cfg2 --only build_bf16 --reps 6 > gpurun_out/c8/ncu1.log 2>&1
timeout 300 python tools/kbench.py cfg2 --only otf --reps 6 > gpurun_out/c8/plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lookup_tc -s 2 -c 1 -o gpurun_out/c8/ltc python tools/kbench.py cfg2 --only otf --reps 6 > gpurun_out/c8/ncu2.log 2>&1
tail -2 gpurun_out/c8/ncu1.log gpurun_out/c8/ncu2.log
