#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python scripts/r02_ens_list.py 2 12 > gpurun_out/r02_run15_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_rb_stream|k_blur_stream" -s 78 -c 4 -o /tmp/ens_kernels python scripts/r02_ens_list.py 2 4 > gpurun_out/r02_run15_ncu.log 2>&1
ncu -i /tmp/ens_kernels.ncu-rep --page raw --csv > gpurun_out/r02_run15_ens_raw.csv 2>/dev/null
ncu -i /tmp/ens_kernels.ncu-rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/r02_run15_ens_source.csv.gz
ls -la /tmp/ens_kernels.ncu-rep gpurun_out/
cat gpurun_out/r02_run15_plain.log | tail -2
