"""Developer tool: one-screen summary of an .ncu-rep (raw page) -- the metrics DESIGN.md / profiles/ quote."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__pcsamp_sample_count", "lts__t_bytes.sum", "l1tex__t_bytes.sum", "launch__grid_size", "launch__block_size",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"]
for path in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        print("==", path, d.get("Kernel Name"))
        for h, u in zip(hdr, units):
            if h in WANT or ("pcsamp_warps_issue_stalled" in h and not h.endswith("not_issued") and float(d[h] or 0) > 0.02 * float(d.get("smsp__pcsamp_sample_count") or 1)):
                print("  %-75s %12s %s" % (h, d[h], u))
