"""Print the metrics we track from an .ncu-rep (run in the build container:
`python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-index]`)."""
import csv, io, subprocess, sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.sum", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.max",
    "lts__t_bytes.sum", "l1tex__t_bytes.sum", "smsp__warps_eligible.avg.per_cycle_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.sum",
    "sm__inst_executed_pipe_fma.sum",
]

def main():
    rep = sys.argv[1]
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2 + which]
    col = {h: i for i, h in enumerate(hdr)}
    print("kernel:", vals[col["Kernel Name"]][:100])
    for k in KEYS:
        if k in col:
            print(f"{k:75s} {vals[col[k]]:>18s} {units[col[k]]}")
    print("-- warp stall reasons (smsp__average_warps_issue_stalled_*_per_issue_active) --")
    st = [(float(vals[i] or 0), h) for h, i in col.items() if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h]
    for v, h in sorted(st, reverse=True)[:8]:
        print(f"  {h.split('issue_stalled_')[1].split('_per_issue')[0]:30s} {v:8.2f}")

main()
