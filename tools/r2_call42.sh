#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c42*
timeout 300 python -m pytest tests/test_gpu_otf_tc.py tests/test_gpu_next.py -m gpu -x -q > gpurun_out/c42_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c42_pytest.log
tail -n 25 gpurun_out/c42_pytest.log
