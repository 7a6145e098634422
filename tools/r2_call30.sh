#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c30*
for b in 1 0; do
  CF_LOOKUP_BARREL=$b timeout 1200 python -m pytest tests/test_gpu_corr.py tests/test_gpu_aten_cuda.py tests/test_gpu_guard_words.py -m gpu -x -q 2>&1 | tail -n 2 >> gpurun_out/c30.log
  for c in cfg2 cfg4; do
    echo "== barrel=$b $c" >> gpurun_out/c30.log
    CF_LOOKUP_BARREL=$b timeout 600 python tools/kbench.py $c --graph --only lookup 2>&1 | grep -i "lookup" >> gpurun_out/c30.log
  done
done
cat gpurun_out/c30.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_kernel -s 20 -c 1 -o gpurun_out/c30_lookup python tools/kbench.py cfg2 --only lookup --reps 3 > gpurun_out/c30_ncu.log 2>&1
python tools/ncu_brief.py gpurun_out/c30_lookup.ncu-rep 2>&1 | head -26
