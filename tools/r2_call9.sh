#!/bin/bash
mkdir -p gpurun_out/c9
timeout 300 python -m pytest tests/test_gpu_otf_tc.py -x -q -m gpu > gpurun_out/c9/pytest_otf.log 2>&1
tail -5 gpurun_out/c9/pytest_otf.log
timeout 300 python tools/kbench.py cfg2 --graph --only otf > gpurun_out/c9/otf_cfg2.log 2>&1
timeout 300 python tools/kbench.py cfg4 --graph --only otf > gpurun_out/c9/otf_cfg4.log 2>&1
cat gpurun_out/c9/otf_cfg2.log gpurun_out/c9/otf_cfg4.log
