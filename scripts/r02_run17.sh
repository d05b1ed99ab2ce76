#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python scripts/r02_ens_list.py 2 4 > gpurun_out/r02_run17_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k "regex:k_rb_stream|k_blur_stream" -s 78 -c 12 -o /tmp/ens_kernels python scripts/r02_ens_list.py 2 4 > gpurun_out/r02_run17_ncu.log 2>&1
ncu -i /tmp/ens_kernels.ncu-rep --page raw --csv > gpurun_out/r02_run17_ens_raw.csv 2>/dev/null
tail -1 gpurun_out/r02_run17_plain.log
