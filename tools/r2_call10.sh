#!/bin/bash
mkdir -p gpurun_out/c10
timeout 300 python -m pytest tests/test_gpu_otf_tc.py tests/test_gpu_build_v2.py -x -q -m gpu > gpurun_out/c10/pytest.log 2>&1
tail -5 gpurun_out/c10/pytest.log
timeout 300 python tools/kbench.py cfg2 --graph --only otf > gpurun_out/c10/otf_cfg2.log 2>&1
timeout 300 python tools/kbench.py cfg4 --graph --only otf > gpurun_out/c10/otf_cfg4.log 2>&1
cat gpurun_out/c10/otf_cfg2.log gpurun_out/c10/otf_cfg4.log
for prec in bf16 bf16x3; do CF_BF16_W32_V1=0 SW_P=29,116 SW_PREC=$prec timeout 300 python tools/build_sweep.py >> gpurun_out/c10/sweep.log 2>&1; done
SW_HW=64 SW_P=64 SW_PREC=bf16 timeout 300 python tools/build_sweep.py >> gpurun_out/c10/sweep.log 2>&1
SW_HW=64 SW_P=64 SW_PREC=bf16x3 timeout 300 python tools/build_sweep.py >> gpurun_out/c10/sweep.log 2>&1
grep build_sweep gpurun_out/c10/sweep.log
