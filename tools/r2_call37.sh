#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c37*
for t in 3 0; do
CF_GEMM_TILED=$t timeout 900 ncu --set full --clock-control none --import-source on -k regex:lookup_kernel -s 20 -c 1 -o gpurun_out/c37_lookup_t$t python tools/kbench.py cfg2 --only lookup --reps 3 > gpurun_out/c37_ncu_t$t.log 2>&1
ncu -i gpurun_out/c37_lookup_t$t.ncu-rep --page raw --csv > gpurun_out/c37_raw_t$t.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/c37_raw_t$t.csv')))
h=rows[0]; v=rows[2]
d=dict(zip(h,v))
print("== tiled=$t duration", d.get('gpu__time_duration.sum'))
for k in sorted(d):
    if k.startswith('l1tex__') and any(s in k for s in ('wavefronts','requests','sectors','data_pipe','t_set','lsuin','throughput','cycles_active')) and d[k] not in ('','0'):
        print(f"  {k:90s} {d[k]}")
PY
done
