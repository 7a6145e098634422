#!/bin/bash
mkdir -p gpurun_out/c17
timeout 300 python tools/kbench.py cfg4 --only otf --reps 6 > gpurun_out/c17/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_tc -s 2 -c 1 -o gpurun_out/c17/ltc python tools/kbench.py cfg4 --only otf --reps 6 > gpurun_out/c17/ncu2.log 2>&1
tail -n 2 gpurun_out/c17/ncu2.log
