#!/bin/bash
mkdir -p gpurun_out/c12
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c12/pytest_all.log 2>&1
tail -6 gpurun_out/c12/pytest_all.log
timeout 300 python bench.py > gpurun_out/c12/bench_cfg2.json 2> gpurun_out/c12/bench_cfg2.err
timeout 300 python bench.py --on-the-fly --no-cpu-baseline --steps 50 > gpurun_out/c12/bench_cfg2_otf.json 2> gpurun_out/c12/bench_cfg2_otf.err
timeout 300 python bench.py --workload cfg4 --no-cpu-baseline --steps 20 > gpurun_out/c12/bench_cfg4.json 2> gpurun_out/c12/bench_cfg4.err
timeout 300 python bench.py --workload cfg4 --on-the-fly --no-cpu-baseline --steps 10 > gpurun_out/c12/bench_cfg4_otf.json 2> gpurun_out/c12/bench_cfg4_otf.err
timeout 300 python bench.py --workload cfg5 --patients 20 > gpurun_out/c12/bench_cfg5.json 2> gpurun_out/c12/bench_cfg5.err
for f in cfg2 cfg2_otf cfg4 cfg4_otf cfg5; do echo "== $f"; cat gpurun_out/c12/bench_$f.json; tail -n 3 gpurun_out/c12/bench_$f.err; done
