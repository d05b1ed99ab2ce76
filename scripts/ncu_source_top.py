"""Top lines of an `ncu --page source --csv` export by stall samples: python scripts/ncu_source_top.py file.csv [lines]"""
import csv
import sys

rows, hdr = [], None
for row in csv.reader(open(sys.argv[1])):
    if not row:
        continue
    if row[0] == "Kernel Name":
        if rows:
            break
        print(row[1][:150])
        continue
    if row[0] == "Address":
        hdr = row
        continue
    if hdr:
        rows.append(row)
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
i_s, i_x = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
total = sum(int(r[i_s]) for r in rows)
print("instructions:", len(rows), "samples:", total)
order = sorted(range(len(rows)), key=lambda k: -int(rows[k][i_s]))[:top]
for k in sorted(order):
    r = rows[k]
    st = sorted(((int(r[c]), hdr[c][6:]) for c in stall_cols if int(r[c])), reverse=True)[:3]
    print("%5d %6.2f%% exec %9s  %-60s %s" % (k, 100.0 * int(r[i_s]) / max(total, 1), r[i_x], r[1].strip()[:60], " ".join("%s:%d" % (n, v) for v, n in st)))
