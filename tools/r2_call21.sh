#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c21.log
timeout 600 python -m pytest tests/test_gpu_warp.py -m gpu -x -q 2>&1 | tail -n 2 >> gpurun_out/c21.log
for m in 0 1; do for pf in 0 16 24; do for t in 4900 9800; do
  echo "== minb8=$m pf=$pf target=$t" >> gpurun_out/c21.log
  CF_WARP_CTAS=$t CF_WARP_MINB8=$m CF_WARP_PF=$pf timeout 300 python tools/kbench.py cfg3 --graph --only warp 2>&1 | grep -E "warp 300x.*smooth" >> gpurun_out/c21.log
done; done; done
cat gpurun_out/c21.log
