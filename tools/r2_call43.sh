#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/c43*
KB_SKIP_SIMT_BWD=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:lookup_bwd -s 2 -c 1 -o gpurun_out/c43_lbwd python tools/kbench.py cfg2 --only bwd --reps 3 > gpurun_out/c43_ncu.log 2>&1
python tools/ncu_brief.py gpurun_out/c43_lbwd.ncu-rep 2>&1 | head -30
ncu -i gpurun_out/c43_lbwd.ncu-rep --page source --csv > gpurun_out/c43_src.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/c43_src.csv')))
hi=[i for i,r in enumerate(rows) if 'Source' in r][0]
h=rows[hi]; ix={c:i for i,c in enumerate(h)}
data=rows[hi+1:]
def f(r,c):
    try: return float(r[ix[c]]) if r[ix[c]] not in ('','-') else 0.0
    except: return 0.0
ts=sum(f(r,'# Samples') for r in data)
top=sorted(range(len(data)), key=lambda i:-f(data[i],'# Samples'))[:25]
for i in sorted(top):
    r=data[i]; print(i, f"{100*f(r,'# Samples')/ts:5.1f}%", r[ix['Source']].strip()[:70], "| lsb", r[ix['stall_long_sb']], "ssb", r[ix['stall_short_sb']], "lg", r[ix['stall_lg']], "mio", r[ix['stall_mio']], "wait", r[ix['stall_wait']])
PY
