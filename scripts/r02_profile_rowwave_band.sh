#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
RW_MEMBERS=148 timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_gs_rowwave" -s 2 -c 1 -o /tmp/rw python scripts/r02_rowwave_band.py 8192 30 1 > gpurun_out/r02_rowwave_band_ncu.log 2>&1
ncu -i /tmp/rw.ncu-rep --page source --csv > /tmp/rw_source.csv 2>/dev/null
python - <<'PY' > gpurun_out/r02_rowwave_band_source_worker.txt 2>&1
import csv
rows, hdr = [], None
for row in csv.reader(open('/tmp/rw_source.csv')):
    if not row: continue
    if row[0] == "Kernel Name": continue
    if row[0] == "Address": hdr = row; continue
    if hdr: rows.append(row)
i_s, i_x = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[i_s]) for r in rows)
wrows = [(k, r) for k, r in enumerate(rows) if k >= 130]
wtot = sum(int(r[i_s]) for k, r in wrows)
print("total samples", tot, "worker (lines >= 130)", wtot)
agg = {}
for k, r in wrows:
    for c in stall_cols:
        agg[hdr[c]] = agg.get(hdr[c], 0) + int(r[c])
print("worker stall totals:", sorted(agg.items(), key=lambda kv: -kv[1])[:10])
order = sorted(wrows, key=lambda kr: -int(kr[1][i_s]))[:110]
for k, r in sorted(order):
    st = sorted(((int(r[c]), hdr[c][6:]) for c in stall_cols if int(r[c])), reverse=True)[:3]
    print("%5d %6.2f%% exec %8s  %-64s %s" % (k, 100.0 * int(r[i_s]) / max(wtot, 1), r[i_x], r[1].strip()[:64], " ".join("%s:%d" % (n, v) for v, n in st)))
PY
cat gpurun_out/r02_rowwave_band_source_worker.txt
