"""Condenses an ncu report (read here with `ncu -i`) into the per-launch metrics DESIGN.md and bench.py cite.
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.csv"""
import csv, io, subprocess, sys
KEEP = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
w = csv.writer(sys.stdout)
stall = [h for h in hdr if "issue_stalled" in h and h.endswith("_per_warp_active.pct")]
cols = [h for h in KEEP if h in hdr] + stall
w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
for h in cols:
    i = hdr.index(h)
    w.writerow([h.replace("smsp__average_warp_latency_issue_stalled_", "stall_").replace("smsp__average_warps_issue_stalled_", "stall_"), units[i]] + [r[i] for r in data])
