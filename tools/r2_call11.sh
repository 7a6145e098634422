#!/bin/bash
mkdir -p gpurun_out/c11
timeout 300 python -m pytest tests/test_gpu_otf_tc.py tests/test_gpu_build_v2.py -x -q -m gpu > gpurun_out/c11/pytest.log 2>&1
tail -3 gpurun_out/c11/pytest.log
timeout 300 python tools/kbench.py cfg4 --graph --only otf > gpurun_out/c11/otf_cfg4.log 2>&1
cat gpurun_out/c11/otf_cfg4.log
for sk in 7 0; do CF_TC2_SKIP=$sk CF_BF16_W32_V1=0 SW_P=29,116 SW_PREC=bf16 timeout 300 python tools/build_sweep.py >> gpurun_out/c11/sweep.log 2>&1; done
grep build_sweep gpurun_out/c11/sweep.log
