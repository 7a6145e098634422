"""Prints a compact summary of an .ncu-rep (development tool): duration, occupancy, issue, DRAM, stall reasons, hot source lines."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
for vals in rows[2:]:
    d = dict(zip(hdr, vals))
    print("kernel:", d.get("Kernel Name", "")[:80])
    keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors_op_read.sum",
            "lts__t_sectors_op_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
            "smsp__average_warp_latency_per_inst_issued.ratio", "lts__t_sector_hit_rate.pct",
            "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
    for k in keys:
        if k in d:
            print(f"  {k:75s} {d[k]}")
    st = sorted(((float(v), k) for k, v in d.items() if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and v), reverse=True)
    for v, k in st[:8]:
        print(f"  stall {k.split('issue_stalled_')[1].split('_per_issue')[0]:25s} {v:.3f}")
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = rows[0] if rows else []
    try:
        si = h.index("Source"); ii = h.index("# Instructions Executed") if "# Instructions Executed" in h else None
        wi = h.index("Warp Stall Sampling (All Samples)") if "Warp Stall Sampling (All Samples)" in h else None
    except ValueError:
        print(h); sys.exit()
    tot = sorted(((float(r[wi] or 0), r[si][:110], r[ii] if ii is not None else "") for r in rows[1:] if len(r) > max(si, wi)), reverse=True)
    for v, s, n in tot[:int(sys.argv[2])]:
        print(f"  {v:8.0f} samp  {n:>10s} inst  {s}")
