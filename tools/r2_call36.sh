#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c36.log
for k in 1 0; do
  CF_GEMM_PDL=$k timeout 900 python -m pytest tests/test_gpu_corr.py -m gpu -x -q 2>&1 | tail -n 1 >> gpurun_out/c36.log
  echo "== pdl=$k" >> gpurun_out/c36.log
  CF_GEMM_PDL=$k timeout 300 python tools/kbench.py cfg2 --graph --only build 2>&1 | grep -E "corr_build\[bf16\]" >> gpurun_out/c36.log
  CF_GEMM_PDL=$k timeout 400 python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step'])" >> gpurun_out/c36.log
  CF_GEMM_PDL=$k timeout 400 python bench.py --no-cpu-baseline --lanes 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench lanes1', d['value'], d['ms_per_step'])" >> gpurun_out/c36.log
done
cat gpurun_out/c36.log
