"""Per-source-line instruction / stall-sample totals from an .ncu-rep captured with --import-source on.
    python tools/ncu_lines.py rep.ncu-rep [kernel-id] [top-N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
kid = sys.argv[2] if len(sys.argv) > 2 else "0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# the output has one block per (kernel launch, file); keep blocks of the first launch only
blocks, cur = [], None
for r in rows:
    if r and r[0] == "File Path":
        cur = {"file": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
seen, out, total_inst, total_samp = set(), [], 0, 0
for b in blocks:
    hdr = next((r for r in b["rows"] if r and r[0] == "Line No"), None)
    if hdr is None:
        continue
    fn = next((r[1] for r in b["rows"] if r and r[0] == "Function Name"), "")
    key = (b["file"], fn)
    if key in seen:
        continue            # a later launch of the same kernel
    seen.add(key)
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    for r in b["rows"]:
        if len(r) > ii and r[0].isdigit():
            num = lambda v: int(v) if v.strip().isdigit() else 0
            inst, samp = num(r[ii]), num(r[si])
            total_inst += inst; total_samp += samp
            out.append((inst, samp, b["file"].split("/")[-1], r[0], r[1].strip()[:110]))
out.sort(reverse=True)
print(f"total inst {total_inst}  samples {total_samp}")
for inst, samp, f, ln, src in out[:top]:
    print(f"{inst:10d} {100*inst/max(total_inst,1):5.1f}%  samp {100*samp/max(total_samp,1):5.1f}%  {f}:{ln}  {src}")
