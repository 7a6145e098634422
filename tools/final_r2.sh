#!/bin/bash
# round 2 evidence run (1 GPU): tests, smoke, every bench workload, launch list, ncu captures.  Outputs -> gpurun_out/r2/
O=gpurun_out/r2
mkdir -p $O
timeout 1800 python -m pytest tests -q -m gpu > $O/r2_pytest_gpu.log 2>&1; tail -n 3 $O/r2_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/r2_smoke.log 2>&1; tail -n 2 $O/r2_smoke.log
timeout 400 python bench.py > $O/r2_bench_cfg2.json 2> $O/err_cfg2.log
timeout 400 python bench.py --precision auto --no-cpu-baseline > $O/r2_bench_cfg2_bf16x3.json 2> $O/err_cfg2x3.log
timeout 400 python bench.py --lanes 1 --no-cpu-baseline > $O/r2_bench_cfg2_lanes1.json 2> $O/err_cfg2l1.log
timeout 400 python bench.py --on-the-fly --no-cpu-baseline --steps 50 > $O/r2_bench_cfg2_on_the_fly.json 2> $O/err_cfg2otf.log
timeout 400 python bench.py --workload cfg1 --no-cpu-baseline > $O/r2_bench_cfg1.json 2> $O/err_cfg1.log
timeout 400 python bench.py --workload cfg3 > $O/r2_bench_cfg3.json 2> $O/err_cfg3.log
timeout 400 python bench.py --workload cfg4 --no-cpu-baseline --steps 20 > $O/r2_bench_cfg4.json 2> $O/err_cfg4.log
timeout 400 python bench.py --workload cfg4 --precision auto --no-cpu-baseline --steps 20 > $O/r2_bench_cfg4_bf16x3.json 2> $O/err_cfg4x3.log
timeout 400 python bench.py --workload cfg4 --on-the-fly --no-cpu-baseline --steps 10 > $O/r2_bench_cfg4_on_the_fly.json 2> $O/err_cfg4otf.log
timeout 600 python bench.py --workload cfg5 > $O/r2_bench_cfg5_n1.json 2> $O/err_cfg5.log
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2_bench_reference.json 2> $O/err_ref.log
for f in $O/r2_bench_*.json; do echo "== $f"; cut -c1-400 $f; done
for c in cfg2 cfg4 cfg3; do timeout 300 python tools/kbench.py $c --graph > $O/r2_kbench_$c.log 2>&1; done
cat $O/r2_kbench_cfg2.log $O/r2_kbench_cfg3.log $O/r2_kbench_cfg4.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline > $O/plain_launch.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 400 --csv --log-file $O/r2_launches_cfg2.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline > $O/ncu_launch.log 2>&1
timeout 300 python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --lanes 1 > $O/plain_l1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_kernel -s 40 -c 2 -o $O/r2_ncu_lookup python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --lanes 1 > $O/ncu_lookup.log 2>&1
timeout 300 python tools/kbench.py cfg2 --only build_bf16x3 --reps 6 > $O/plain_b.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_tc2 -s 2 -c 1 -o $O/r2_ncu_corr_tc2 python tools/kbench.py cfg2 --only build_bf16x3 --reps 6 > $O/ncu_b.log 2>&1
timeout 300 python tools/kbench.py cfg4 --only otf --reps 6 > $O/plain_o.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_tc -s 2 -c 1 -o $O/r2_ncu_lookup_tc python tools/kbench.py cfg4 --only otf --reps 6 > $O/ncu_o.log 2>&1
timeout 300 python tools/kbench.py cfg4 --only lookup --reps 6 > $O/plain_l4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:lookup_kernel -s 30 -c 1 -o $O/r2_ncu_lookup_cfg4 python tools/kbench.py cfg4 --only lookup --reps 6 > $O/ncu_l4.log 2>&1
timeout 300 python tools/kbench.py cfg3 --only warp --reps 6 > $O/plain_w.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:warp2d_tile -s 6 -c 2 -o $O/r2_ncu_warp python tools/kbench.py cfg3 --only warp --reps 6 > $O/ncu_w.log 2>&1
KB_SKIP_SIMT_BWD=1 timeout 300 python tools/kbench.py cfg2 --only bwd --reps 6 > $O/plain_bw.log 2>&1 &&
KB_SKIP_SIMT_BWD=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_bwd_tc -s 2 -c 1 -o $O/r2_ncu_corr_bwd_tc python tools/kbench.py cfg2 --only bwd --reps 6 > $O/ncu_bw.log 2>&1
for r in lookup corr_tc2 lookup_tc lookup_cfg4 warp corr_bwd_tc; do python tools/ncu_brief.py $O/r2_ncu_$r.ncu-rep > $O/r2_ncu_$r.txt 2>&1; done
ls -la $O
