"""Summarise an .ncu-rep (here, no GPU needed): the metrics the profiles/ summaries quote, one column per launch.
usage: python scripts/ncu_summary.py report.ncu-rep [kernel-name-substring]"""
import csv, io, subprocess, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_static", "sm__maximum_warps_per_active_cycle_pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "smsp__cycles_active.avg",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
if len(sys.argv) > 2:
    data = [r for r in data if sys.argv[2] in r[hdr.index("Kernel Name")]]
print("launches:", [r[hdr.index("Kernel Name")][:60] for r in data])
cols = [c for c in hdr if c in WANT or "issue_stalled" in c and c.endswith("per_issue_active.ratio")]
for c in cols:
    i = hdr.index(c)
    print("%-90s %-16s %s" % (c, units[i], "  ".join(r[i] for r in data)))
