#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c35.log
for pk in 1 0; do
  CF_WARP_PACK=$pk timeout 900 python -m pytest tests/test_gpu_warp.py tests/test_gpu_aten_cuda.py tests/test_gpu_guard_words.py -m gpu -x -q 2>&1 | tail -n 1 >> gpurun_out/c35.log
  echo "== pack=$pk" >> gpurun_out/c35.log
  CF_WARP_PACK=$pk timeout 300 python tools/kbench.py cfg3 --graph --only warp 2>&1 | grep -E "warp 300x" >> gpurun_out/c35.log
  CF_WARP_PACK=$pk timeout 300 python tools/kbench.py cfg2 --graph --only warp 2>&1 | grep -E "^warp " >> gpurun_out/c35.log
done
cat gpurun_out/c35.log
