#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, on the CPU box): per-launch key metrics and the hottest SASS lines."""
import csv, subprocess, sys, io
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__cycles_active.avg", "sm__cycles_active.max",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__t_bytes.sum"]
for r in rows[2:]:
    print("===", r[hdr.index("ID")], r[hdr.index("Kernel Name")][:60])
    for w in want:
        if w in hdr:
            print("   %-62s %s %s" % (w, r[hdr.index(w)], units[hdr.index(w)]))
    stalls = [(float(r[i] or 0), h) for i, h in enumerate(hdr) if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    print("   stalls:", ", ".join("%s=%.2f" % (h.split("issue_stalled_")[1].split("_per_")[0], v) for v, h in sorted(stalls, reverse=True)[:6]))
