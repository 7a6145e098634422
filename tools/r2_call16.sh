#!/bin/bash
# warp grid sweep (direct and staged)
mkdir -p gpurun_out
for st in 0 1; do for t in 588 882 1176 1764 2368 2646 3528 4900; do
  echo "== staged=$st target=$t" >> gpurun_out/c16.log
  CF_WARP_STAGED=$st CF_WARP_CTAS=$t timeout 300 python tools/kbench.py cfg3 --graph --only warp 2>&1 | grep -E "warp 300x(1|5)x.*smooth" >> gpurun_out/c16.log
done; done
cat gpurun_out/c16.log
