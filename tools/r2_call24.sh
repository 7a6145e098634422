#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c24*
timeout 900 python -m pytest tests/test_gpu_bwd_tc.py -m gpu -x -q > gpurun_out/c24_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c24_pytest.log
tail -n 3 gpurun_out/c24_pytest.log
KB_SKIP_SIMT_BWD=1 timeout 600 python tools/kbench.py cfg2 --graph --only bwd > gpurun_out/c24_kb_cfg2.log 2>&1; grep -i "volume backward" gpurun_out/c24_kb_cfg2.log
KB_SKIP_SIMT_BWD=1 timeout 600 python tools/kbench.py cfg4 --graph --only bwd > gpurun_out/c24_kb_cfg4.log 2>&1; grep -i "volume backward" gpurun_out/c24_kb_cfg4.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_bwd_tc -c 2 -o gpurun_out/c24_bwd_tc python tools/kbench.py cfg2 --only bwd --reps 2 > gpurun_out/c24_ncu.log 2>&1; tail -n 2 gpurun_out/c24_ncu.log
python tools/ncu_brief.py gpurun_out/c24_bwd_tc.ncu-rep > gpurun_out/c24_ncu_brief.txt 2>&1; head -70 gpurun_out/c24_ncu_brief.txt
