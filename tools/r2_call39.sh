#!/bin/bash
# final confirmation on the committed tree; every step bounded
mkdir -p gpurun_out/r2
O=gpurun_out/r2
timeout 600 python -m pytest tests -q -x -m gpu > $O/r2_pytest_gpu.log 2>&1; echo "rc=$?" >> $O/r2_pytest_gpu.log; tail -n 3 $O/r2_pytest_gpu.log
timeout 200 python __graft_entry__.py --smoke > $O/r2_smoke.log 2>&1; tail -n 1 $O/r2_smoke.log
timeout 300 python bench.py > $O/r2_bench_cfg2.json 2> $O/err_cfg2.log; cut -c1-200 $O/r2_bench_cfg2.json
timeout 300 python bench.py --workload cfg3 --no-cpu-baseline > $O/r2_bench_cfg3.json 2> $O/err_cfg3.log; cut -c1-200 $O/r2_bench_cfg3.json
timeout 300 python bench.py --workload cfg4 --no-cpu-baseline --steps 20 > $O/r2_bench_cfg4.json 2> $O/err_cfg4.log; cut -c1-200 $O/r2_bench_cfg4.json
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference.json 2> $O/err_ref.log; cut -c1-160 $O/r2_bench_reference.json
