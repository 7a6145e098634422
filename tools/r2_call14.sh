#!/bin/bash
mkdir -p gpurun_out/c14
timeout 900 python -m pytest tests/test_gpu_aten_cuda.py tests/test_gpu_warp.py tests/test_gpu_corr.py -q -m gpu > gpurun_out/c14/pytest.log 2>&1
tail -25 gpurun_out/c14/pytest.log
timeout 300 python tools/kbench.py cfg2 --graph --only bwd > gpurun_out/c14/bwd_cfg2.log 2>&1
cat gpurun_out/c14/bwd_cfg2.log
timeout 300 python tools/kbench.py cfg3 --graph --only warp > gpurun_out/c14/warp_cfg3.log 2>&1
cat gpurun_out/c14/warp_cfg3.log
