#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=300 -x -k "row_pipeline or wavefront or golden or reference or host_demo" -rf > gpurun_out/r02_rwi_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_rwi_tests.log
tail -3 gpurun_out/r02_rwi_tests.log
timeout 300 python scripts/r02_rowwave_micro.py 8192 2>&1 | head -6
timeout 900 python scripts/r02_probe3.py wavefront_rows > gpurun_out/r02_rwi_wavefront_modes.jsonl 2>&1
cut -c1-100 gpurun_out/r02_rwi_wavefront_modes.jsonl
