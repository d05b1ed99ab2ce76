"""Summarise `ncu --page raw --csv` exports (one file per capture) into a small table: duration, instructions,
issue utilisation, DRAM bytes, L1/L2 hit rates and the top stall reasons.  Usage: ncu_summary.py file.raw.csv ..."""
import csv
import sys

KEYS = [("gpu__time_duration.sum", "duration_us"), ("smsp__inst_executed.sum", "warp_instructions"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
        ("launch__registers_per_thread", "registers"), ("dram__bytes_read.sum", "dram_read_MB"),
        ("dram__bytes_write.sum", "dram_write_MB"), ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"),
        ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
        ("gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "mem_throughput_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct")]

for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    if len(rows) < 3:
        print("## %s: empty" % path)
        continue
    h, units = rows[0], rows[1]
    for v in rows[2:]:
        print("## %s  [%s]" % (v[h.index("Kernel Name")][:90], path.split("/")[-1]))
        for k, name in KEYS:
            if k in h:
                val, u = v[h.index(k)], units[h.index(k)]
                print("   %-22s %s %s" % (name, val, u if name in ("duration_us", "dram_read_MB", "dram_write_MB") else ""))
        st = sorted(((float(v[i]), x.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
                     for i, x in enumerate(h) if "issue_stalled" in x and x.endswith("per_issue_active.ratio")), reverse=True)
        print("   stalls (warps per issue): " + ", ".join("%s %.2f" % (b, a) for a, b in st[:6]))
