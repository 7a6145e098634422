#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c45*
timeout 300 python -m pytest tests/test_gpu_next.py tests/test_gpu_warp.py tests/test_gpu_autograd_r2.py -m gpu -x -q > gpurun_out/c45_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c45_pytest.log
tail -n 4 gpurun_out/c45_pytest.log
KB_SKIP_SIMT_BWD=1 timeout 200 python tools/kbench.py cfg2 --graph --only bwd 2>&1 | grep -i "warp backward" | tee gpurun_out/c45_kb.log
KB_SKIP_SIMT_BWD=1 timeout 200 python tools/kbench.py cfg4 --graph --only bwd 2>&1 | grep -i "warp backward" | tee -a gpurun_out/c45_kb.log
