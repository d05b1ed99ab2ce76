#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=300 -k "wavefront" -rf > gpurun_out/r02_run3_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run3_tests.log
timeout 900 python scripts/r02_probe3.py all > gpurun_out/r02_run3_probe.jsonl 2> gpurun_out/r02_run3_probe.err
echo "probe rc=$?" >> gpurun_out/r02_run3_probe.err
timeout 300 python scripts/r02_advect_probe.py 8192 1 --quick > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_advect_rows -s 2 -c 2 -o gpurun_out/r02_advect_rows_v2 python scripts/r02_advect_probe.py 8192 1 --quick > gpurun_out/r02_run3_ncu.log 2>&1
tail -5 gpurun_out/r02_run3_tests.log
cat gpurun_out/r02_run3_probe.jsonl
