#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c29*
timeout 1200 python -m pytest tests/test_gpu_next.py tests/test_gpu_autograd_r2.py tests/test_gpu_bwd_tc.py -m gpu -x -q > gpurun_out/c29_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c29_pytest.log
tail -n 4 gpurun_out/c29_pytest.log
KB_SKIP_SIMT_BWD=1 timeout 600 python tools/kbench.py cfg2 --graph --only bwd > gpurun_out/c29_kb_cfg2.log 2>&1; grep -i "backward" gpurun_out/c29_kb_cfg2.log
KB_SKIP_SIMT_BWD=1 timeout 600 python tools/kbench.py cfg4 --graph --only bwd > gpurun_out/c29_kb_cfg4.log 2>&1; grep -i "backward" gpurun_out/c29_kb_cfg4.log
