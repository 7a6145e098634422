"""Per-source-line instruction totals of the N-th profiled launch of a kernel matching PATTERN.
    python tools/ncu_fn_lines.py rep.ncu-rep PATTERN [occurrence=0] [min_pct=0.4]"""
import csv, io, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
minpct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.4
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "File Path":
        cur = {"file": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
count = collections.Counter()
tot, out = 0, []
for b in blocks:
    fn = next((r[1] for r in b["rows"] if r and r[0] == "Function Name"), "")
    hdr = next((r for r in b["rows"] if r and r[0] == "Line No"), None)
    if hdr is None or pat not in fn:
        continue
    key = (b["file"], fn)
    k = count[key]
    count[key] += 1
    if k != occ:
        continue
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    for r in b["rows"]:
        if len(r) > ii and r[0].isdigit():
            inst = int(r[ii]) if r[ii].strip().isdigit() else 0
            s = int(r[si]) if r[si].strip().isdigit() else 0
            tot += inst
            out.append((b["file"].split("/")[-1], int(r[0]), inst, s, r[1].strip()[:120]))
print("total warp-instructions", tot)
for f, ln, inst, s, src in out:
    if inst > tot * minpct / 100:
        print(f"{f}:{ln:4d} {inst:9d} {100 * inst / tot:5.1f}%  samp {s:5d}  {src}")
