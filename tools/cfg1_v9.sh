# cfg1 (BASELINE configs[0], the reference's own CPU-runnable case) on the GPU box: this build, the CPU arm with all host threads and
# with one thread (per-core figure), and the reference "as actually run" with torch CUDA eager ops.
mkdir -p gpurun_out/v9; O=gpurun_out/v9
python bench.py --workload cfg1 > $O/r1_bench_cfg1_v9.json 2>$O/cfg1.err; tail -2 $O/cfg1.err
python bench.py --impl reference --workload cfg1 --steps 20 --warmup 2 > $O/r1_bench_ref_cfg1_v9.json 2>/dev/null
python bench.py --impl reference --workload cfg1 --steps 20 --warmup 2 --cpu-threads 1 > $O/r1_bench_ref_cfg1_1thread_v9.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 --cpu-threads 1 > $O/r1_bench_ref_cfg2_1thread_v9.json 2>/dev/null
python bench.py --impl reference-gpu --workload cfg1 --steps 20 > $O/r1_bench_refgpu_cfg1_v9.json 2>/dev/null
for f in r1_bench_cfg1_v9 r1_bench_ref_cfg1_v9 r1_bench_ref_cfg1_1thread_v9 r1_bench_ref_cfg2_1thread_v9 r1_bench_refgpu_cfg1_v9; do
  python -c "
import json; d=json.load(open('$O/$f.json')); print('$f', round(d['value'],1), round(d['ms_per_step'],4), d.get('e2e',{}).get('value'), d.get('cpu_baseline',{}).get('cores'), d.get('batched',{}).get('value'), d.get('roofline',{}).get('frac'))"
done
