#!/bin/bash
mkdir -p gpurun_out/c19
CF_BF16_W32_V1=0 timeout 600 python -m pytest tests/test_gpu_build_v2.py tests/test_gpu_corr.py -x -q -m gpu > gpurun_out/c19/pytest.log 2>&1
tail -n 4 gpurun_out/c19/pytest.log
for nb in 1 2; do
  echo "== NB=$nb" >> gpurun_out/c19/sweep.log
  CF_TC2_NB=$nb CF_BF16_W32_V1=0 SW_P=29,116 SW_PREC=bf16 timeout 300 python tools/build_sweep.py >> gpurun_out/c19/sweep.log 2>&1
  CF_TC2_NB=$nb SW_HW=64 SW_P=64 SW_PREC=bf16 timeout 300 python tools/build_sweep.py >> gpurun_out/c19/sweep.log 2>&1
done
echo "== v1" >> gpurun_out/c19/sweep.log
SW_P=29,116 SW_PREC=bf16 timeout 300 python tools/build_sweep.py >> gpurun_out/c19/sweep.log 2>&1
grep -E "==|build_sweep" gpurun_out/c19/sweep.log
