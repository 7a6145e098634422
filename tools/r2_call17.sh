#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c17.log
for t in 4900 7350 9800 14700 29400; do
  echo "== staged=0 target=$t" >> gpurun_out/c17.log
  CF_WARP_STAGED=0 CF_WARP_CTAS=$t timeout 300 python tools/kbench.py cfg3 --graph --only warp 2>&1 | grep -E "warp 300x" >> gpurun_out/c17.log
done
cat gpurun_out/c17.log
