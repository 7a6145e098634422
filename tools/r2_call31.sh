#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c31*
timeout 1200 python -m pytest tests/test_gpu_otf_tc.py tests/test_gpu_next.py -m gpu -x -q 2>&1 | tail -n 3 > gpurun_out/c31.log
for c in cfg2 cfg4; do timeout 600 python tools/kbench.py $c --graph --only otf 2>&1 | grep -i "on-the-fly" >> gpurun_out/c31.log; done
timeout 400 python bench.py --workload cfg4 --on-the-fly --no-cpu-baseline --steps 10 > gpurun_out/c31_bench_cfg4_otf.json 2> gpurun_out/c31_err.log
cut -c1-300 gpurun_out/c31_bench_cfg4_otf.json >> gpurun_out/c31.log
cat gpurun_out/c31.log
