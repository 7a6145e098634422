#!/bin/bash
mkdir -p gpurun_out/c7
timeout 600 python -m pytest tests/test_gpu_build_v2.py tests/test_gpu_corr.py -q -m gpu -x > gpurun_out/c7/pytest.log 2>&1
tail -5 gpurun_out/c7/pytest.log
for prec in bf16 bf16x3; do SW_PREC=$prec timeout 300 python tools/build_sweep.py >> gpurun_out/c7/sweep.log 2>&1; done
SW_HW=64 SW_P=16,64 SW_PREC=bf16 timeout 300 python tools/build_sweep.py >> gpurun_out/c7/sweep.log 2>&1
SW_HW=64 SW_P=16,64 SW_PREC=bf16 CF_TC2_SCHED=0 timeout 300 python tools/build_sweep.py >> gpurun_out/c7/sweep.log 2>&1
SW_HW=64 SW_P=64 SW_PREC=bf16x3 timeout 300 python tools/build_sweep.py >> gpurun_out/c7/sweep.log 2>&1
grep build_sweep gpurun_out/c7/sweep.log
timeout 300 python tools/kbench.py cfg2 --graph --only otf,lookup > gpurun_out/c7/otf_cfg2.log 2>&1
timeout 300 python tools/kbench.py cfg4 --graph --only otf,lookup > gpurun_out/c7/otf_cfg4.log 2>&1
cat gpurun_out/c7/otf_cfg2.log gpurun_out/c7/otf_cfg4.log
