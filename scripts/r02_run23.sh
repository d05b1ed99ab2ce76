#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for m in 3 0; do timeout 120 python scripts/r02_rowwave_one.py 2048 $m 3; timeout 120 python scripts/r02_rowwave_one.py 2048 $m 3 10.49 42.96; done > gpurun_out/r02_run23_plain.log 2>&1
cat gpurun_out/r02_run23_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_gs_rowwave" -s 1 -c 1 -o /tmp/rw python scripts/r02_rowwave_one.py 2048 3 1 > gpurun_out/r02_run23_ncu.log 2>&1
ncu -i /tmp/rw.ncu-rep --page raw --csv > gpurun_out/r02_run23_raw.csv 2>/dev/null
ncu -i /tmp/rw.ncu-rep --page source --csv > /tmp/rw_source.csv 2>/dev/null
python scripts/ncu_source_top.py /tmp/rw_source.csv 45 > gpurun_out/r02_run23_source_top.txt 2>&1
python scripts/ncu_summary.py gpurun_out/r02_run23_raw.csv
cat gpurun_out/r02_run23_source_top.txt
