#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c18.log
timeout 600 python -m pytest tests/test_gpu_warp.py -m gpu -x -q 2>&1 | tail -n 2 >> gpurun_out/c18.log
for pf in 0 16 24 32; do for t in 882 2450 4900 9800; do
  echo "== pf=$pf target=$t" >> gpurun_out/c18.log
  CF_WARP_PF=$pf CF_WARP_CTAS=$t timeout 300 python tools/kbench.py cfg3 --graph --only warp 2>&1 | grep -E "warp 300x(1|5)x" >> gpurun_out/c18.log
done; done
cat gpurun_out/c18.log
