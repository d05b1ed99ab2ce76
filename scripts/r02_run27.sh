#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=240 -x -k "row_pipeline" -rf > gpurun_out/r02_run27_tests_a.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run27_tests_a.log
tail -5 gpurun_out/r02_run27_tests_a.log
timeout 300 python scripts/r02_rowwave_micro.py 8192 2>&1 | tail -14
for n in 1024 2048 8192; do timeout 120 python scripts/r02_rowwave_one.py $n 3 3; timeout 120 python scripts/r02_rowwave_one.py $n 3 3 10.49 42.96; done 2>&1
