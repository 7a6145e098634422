#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c20.log
timeout 600 python -m pytest tests/test_gpu_warp.py -m gpu -x -q 2>&1 | tail -n 2 >> gpurun_out/c20.log
for a in 0 -1 6 24; do for t in 4900 9800 14700; do
  echo "== minb8=1 pf=16 ahead=$a target=$t" >> gpurun_out/c20.log
  CF_WARP_AHEAD=$a CF_WARP_CTAS=$t CF_WARP_MINB8=1 CF_WARP_PF=16 timeout 300 python tools/kbench.py cfg3 --graph --only warp 2>&1 | grep -E "warp 300x(1|5)x.*smooth" >> gpurun_out/c20.log
done; done
cat gpurun_out/c20.log
