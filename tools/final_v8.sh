set -x
mkdir -p gpurun_out/v8; O=gpurun_out/v8
python -m pytest tests -q -m gpu 2>&1 | tail -4 > $O/r1_pytest_gpu_v8.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r1_smoke_v8.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > $O/r1_bench_ref_v8.json 2>$O/ref.err
python bench.py > $O/r1_bench_cfg2_v8.json 2>$O/cfg2.err
python bench.py --lanes 1 --no-cpu-baseline > $O/r1_bench_cfg2_lanes1_v8.json 2>/dev/null
python bench.py --workload cfg3 > $O/r1_bench_cfg3_v8.json 2>/dev/null
python bench.py --workload cfg4 --steps 30 > $O/r1_bench_cfg4_v8.json 2>/dev/null
python bench.py --workload cfg5 > $O/r1_cohort_n1_v8.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r1_launches_cfg2_v8.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline > $O/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lookup_kernel -s 40 -c 2 -o $O/lookup_v8 python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --lanes 1 > $O/ncu_full.log 2>&1
ls -la $O
