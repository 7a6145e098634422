"""Regenerates profiles/r2/sass_tensor_tma_kernels.txt: per-kernel counts of the tensor-core / TMA / async-copy / packed-fp32 SASS mnemonics
in the objects built by __graft_entry__.build() (cuobjdump -sass; no GPU needed)."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "cardiac-segmentation-optical-flow_b200", "build")
FILES = ["corr_tc2", "lookup_tc", "corr_bwd_tc", "corr_tcgen05", "lookup", "warp", "backward"]
PAT = re.compile(r"\b(UTCHMMA|UTCQMMA|LDTM(?:\.x\d+)?|UTMALDG(?:\.\dD)?|UTMASTG(?:\.\dD)?|UTCBAR|UTCATOMSWS|SYNCS|LDGSTS(?:\.E\.BYPASS\.128)?|UBLKPF|UBLKCP|FFMA2|FMUL2|FADD2|REDUX|SHFL|RED)\b")
out = ["# SASS evidence (round 2): `cuobjdump -sass cardiac-segmentation-optical-flow_b200/build/<file>.o`, mnemonic counts per kernel.",
       "# Built by __graft_entry__.build() with nvcc 12.9, -gencode arch=compute_100a,code=sm_100a.  Regenerate: python tools/sass_counts.py", ""]
for f in FILES:
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, f + ".o")], capture_output=True, text=True).stdout
    out += ["", f"## {f}.cu", ""]
    name, cnt, n = None, collections.Counter(), 0
    def flush():
        if name:
            ms = ", ".join(f"{k} x{v}" for k, v in sorted(cnt.items()))
            out.extend([name[:90], f"    instructions ~{n}; " + (ms if ms else "no tensor/TMA/async mnemonics"), ""])
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            name, cnt, n = m.group(1), collections.Counter(), 0
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", line):
            n += 1
            mm = PAT.search(line.split("*/", 1)[1])
            if mm:
                cnt[mm.group(1)] += 1
    flush()
open(os.path.join(ROOT, "profiles", "r2", "sass_tensor_tma_kernels.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out[:60]))
