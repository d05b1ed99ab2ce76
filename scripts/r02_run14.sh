#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python scripts/r02_ens_list.py 2 12 > gpurun_out/r02_run14_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_run14_ens_list_mode2.csv python scripts/r02_ens_list.py 2 12 > gpurun_out/r02_run14_ncu.log 2>&1
cat gpurun_out/r02_run14_plain.log | tail -2
