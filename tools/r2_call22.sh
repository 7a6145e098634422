#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c22*.log
timeout 900 python -m pytest tests/test_gpu_bwd_tc.py -m gpu -x -q -s > gpurun_out/c22_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c22_pytest.log
tail -n 30 gpurun_out/c22_pytest.log
timeout 600 python tools/kbench.py cfg2 --graph --only bwd > gpurun_out/c22_kb_cfg2.log 2>&1; grep -i "backward" gpurun_out/c22_kb_cfg2.log
