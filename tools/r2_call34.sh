#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c34.log
for m in 0 1; do
  CF_LOOKUP_LEVEL_MAJOR=$m timeout 600 python -m pytest tests/test_gpu_corr.py -m gpu -x -q 2>&1 | tail -n 1 >> gpurun_out/c34.log
  for c in cfg2 cfg4; do
    echo "== level_major=$m $c" >> gpurun_out/c34.log
    CF_LOOKUP_LEVEL_MAJOR=$m timeout 600 python tools/kbench.py $c --graph --only lookup 2>&1 | grep -i "lookup" >> gpurun_out/c34.log
  done
  CF_LOOKUP_LEVEL_MAJOR=$m timeout 400 python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('bench', d['value'], r['frac'], r['as_run_in_the_step']['frac'])" >> gpurun_out/c34.log
done
cat gpurun_out/c34.log
