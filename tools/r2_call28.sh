#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c28*
for t in 3 0; do for nw in 3 2 4; do
  echo "== CF_GEMM_TILED=$t NW=$nw" >> gpurun_out/c28.log
  CF_GEMM_TILED=$t CF_LOOKUP_NW=$nw timeout 600 python tools/kbench.py cfg2 --graph --only lookup 2>&1 | grep -i "lookup" >> gpurun_out/c28.log
done; done
for t in 3 0; do
  echo "== cfg4 CF_GEMM_TILED=$t" >> gpurun_out/c28.log
  CF_GEMM_TILED=$t timeout 600 python tools/kbench.py cfg4 --graph --only lookup 2>&1 | grep -i "lookup" >> gpurun_out/c28.log
done
cat gpurun_out/c28.log
