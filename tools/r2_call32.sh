#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c32*
timeout 300 python tools/kbench.py cfg4 --only otf --reps 6 > gpurun_out/c32_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_tc -s 2 -c 1 -o gpurun_out/c32_k4 python tools/kbench.py cfg4 --only otf --reps 6 > gpurun_out/c32_ncu.log 2>&1
python tools/ncu_brief.py gpurun_out/c32_k4.ncu-rep 2>&1 | head -28
