#!/bin/bash
mkdir -p gpurun_out/c13
for st in 0 500 1000 2000 4000; do
  echo "== stagger $st" >> gpurun_out/c13/lk.log
  CF_LOOKUP_STAGGER=$st timeout 200 python tools/kbench.py cfg2 --graph --only lookup >> gpurun_out/c13/lk.log 2>&1
done
grep -E "==|x12" gpurun_out/c13/lk.log
