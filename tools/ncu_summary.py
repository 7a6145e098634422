"""Concise per-kernel summary of an .ncu-rep (run where ncu is installed; no GPU needed).
    python tools/ncu_summary.py gpurun_out/x.ncu-rep [--stalls]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEYS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("lts__t_sector_hit_rate.pct", "l2_hit%"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("launch__registers_per_thread", "regs"),
        ("sm__inst_executed.sum", "inst"), ("smsp__inst_executed.sum", "inst"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1pipe%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wf"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bank_confl"),
        ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "gld_sect"), ("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "gld_req"),
        ("l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct", "l1_hit%"),
        ("lts__t_bytes.sum", "l2_bytes"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu%"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__occupancy_limit_shared_mem", "lim_smem"),
        ("launch__occupancy_limit_registers", "lim_regs"), ("launch__occupancy_limit_warps", "lim_warps")]
STALL = [h for h in hdr if h.startswith("smsp__average_warp") and h.endswith("_per_issue_active.ratio")] if "--stalls" in sys.argv else []
seen = set()
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")][:70]
    out = []
    for k, short in KEYS:
        if k in hdr and r[hdr.index(k)] not in ("", "n/a"):
            v = r[hdr.index(k)].replace(",", "")
            try:
                f = float(v); v = f"{f:.4g}"
            except ValueError:
                pass
            out.append(f"{short}={v}{units[hdr.index(k)] if units[hdr.index(k)] not in ('%','') else ''}")
    print(name); print("   " + "  ".join(out))
    if STALL:
        st = sorted(((float(r[hdr.index(h)] or 0), h.replace("smsp__average_warps_issue_stalled_", "").replace("smsp__average_warp_latency_issue_stalled_","").replace("_per_issue_active.ratio", "")) for h in STALL), reverse=True)[:7]
        print("   stalls: " + "  ".join(f"{n}={v:.2f}" for v, n in st))
