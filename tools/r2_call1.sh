#!/bin/bash
# round 2, GPU call 1: new tests (build v2, autograd holes), then per-kernel timings and the default bench line
set -x
mkdir -p gpurun_out/c1
timeout 900 python -m pytest tests/test_gpu_build_v2.py tests/test_gpu_autograd_r2.py -x -q -m gpu > gpurun_out/c1/pytest_new.log 2>&1
echo "new tests rc=$?" >> gpurun_out/c1/pytest_new.log
tail -15 gpurun_out/c1/pytest_new.log
timeout 1500 python -m pytest tests -q -m gpu -x --deselect tests/test_gpu_build_v2.py --deselect tests/test_gpu_autograd_r2.py > gpurun_out/c1/pytest_all.log 2>&1
echo "all tests rc=$?" >> gpurun_out/c1/pytest_all.log
tail -8 gpurun_out/c1/pytest_all.log
timeout 300 python tools/kbench.py cfg2 --graph --only build_bf16,build_bf16x3,lookup > gpurun_out/c1/kbench_cfg2.log 2>&1
timeout 300 python tools/kbench.py cfg4 --graph --only build_bf16,build_bf16x3,lookup > gpurun_out/c1/kbench_cfg4.log 2>&1
cat gpurun_out/c1/kbench_cfg2.log gpurun_out/c1/kbench_cfg4.log
timeout 300 python bench.py > gpurun_out/c1/bench_cfg2.json 2> gpurun_out/c1/bench_cfg2.err
timeout 300 python bench.py --precision bf16 --no-cpu-baseline > gpurun_out/c1/bench_cfg2_bf16.json 2> gpurun_out/c1/bench_cfg2_bf16.err
timeout 300 python bench.py --workload cfg4 --no-cpu-baseline --steps 20 > gpurun_out/c1/bench_cfg4.json 2> gpurun_out/c1/bench_cfg4.err
cat gpurun_out/c1/bench_cfg2.json gpurun_out/c1/bench_cfg2_bf16.json gpurun_out/c1/bench_cfg4.json
tail -3 gpurun_out/c1/*.err
