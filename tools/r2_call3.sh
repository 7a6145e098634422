#!/bin/bash
set -x
mkdir -p gpurun_out/c3
timeout 600 python -m pytest tests/test_gpu_build_v2.py tests/test_gpu_autograd_r2.py -q -m gpu > gpurun_out/c3/pytest_new.log 2>&1
tail -5 gpurun_out/c3/pytest_new.log
for sk in 0 1 6 7; do
  echo "== CF_TC2_SKIP=$sk" >> gpurun_out/c3/skip.log
  CF_TC2_SKIP=$sk timeout 200 python tools/kbench.py cfg2 --graph --only build_bf16,build_bf16x3 >> gpurun_out/c3/skip.log 2>&1
done
timeout 200 python tools/kbench.py cfg4 --graph --only build_bf16,build_bf16x3 >> gpurun_out/c3/skip.log 2>&1
grep -E "==|corr_build" gpurun_out/c3/skip.log
timeout 300 python -m pytest tests/test_gpu_otf_tc.py -x -q -m gpu > gpurun_out/c3/pytest_otf.log 2>&1
tail -30 gpurun_out/c3/pytest_otf.log
