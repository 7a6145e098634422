#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c40*
timeout 200 python -m pytest tests/test_gpu_bwd_tc.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/c40_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c40_pytest.log
tail -n 3 gpurun_out/c40_pytest.log
KB_SKIP_SIMT_BWD=1 timeout 200 python tools/kbench.py cfg2 --graph --only bwd 2>&1 | grep -i "volume backward" | tee gpurun_out/c40_kb.log
KB_SKIP_SIMT_BWD=1 timeout 200 python tools/kbench.py cfg4 --graph --only bwd 2>&1 | grep -i "volume backward" | tee -a gpurun_out/c40_kb.log
