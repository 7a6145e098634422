# Round-1 state v9 evidence (N = 1): run on the GPU box through gpurun.
set -x
mkdir -p gpurun_out/v9; O=gpurun_out/v9
python -m pytest tests -q -m gpu 2>&1 | tail -4 > $O/r1_pytest_gpu_v9.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r1_smoke_v9.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > $O/r1_bench_ref_v9.json 2>$O/ref.err
python bench.py > $O/r1_bench_cfg2_v9.json 2>$O/cfg2.err
python bench.py --workload cfg3 > $O/r1_bench_cfg3_v9.json 2>/dev/null
python bench.py --workload cfg4 --steps 30 > $O/r1_bench_cfg4_v9.json 2>/dev/null
python bench.py --workload cfg5 > $O/r1_cohort_n1_v9.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r1_launches_cfg2_v9.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline > $O/ncu_launch.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r1_launches_cfg5_unit_v9.csv python bench.py --workload cfg5 --patients 1 --slices 2 --no-graph --lanes 1 > $O/ncu_launch5.log 2>&1
ls -la $O
