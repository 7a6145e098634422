#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c41*
timeout 300 python -m pytest tests/test_gpu_next.py tests/test_gpu_autograd_r2.py tests/test_gpu_bwd_tc.py -m gpu -x -q > gpurun_out/c41_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c41_pytest.log
tail -n 4 gpurun_out/c41_pytest.log
KB_SKIP_SIMT_BWD=1 timeout 200 python tools/kbench.py cfg2 --graph --only bwd 2>&1 | grep -i "lookup backward" | tee gpurun_out/c41_kb.log
KB_SKIP_SIMT_BWD=1 timeout 200 python tools/kbench.py cfg4 --graph --only bwd 2>&1 | grep -i "lookup backward" | tee -a gpurun_out/c41_kb.log
