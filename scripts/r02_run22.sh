#!/bin/bash
# Row pipeline for the reference-order solve: its own tests first (bounded), then every wavefront test, then timings
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=240 -x -k "row_pipeline" -rf > gpurun_out/r02_run22_tests_a.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run22_tests_a.log
tail -15 gpurun_out/r02_run22_tests_a.log
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=300 -k "wavefront or golden or reference or host_demo or bordered" -rf > gpurun_out/r02_run22_tests_b.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run22_tests_b.log
tail -8 gpurun_out/r02_run22_tests_b.log
timeout 600 python scripts/r02_probe3.py wavefront_rows > gpurun_out/r02_run22_wavefront.jsonl 2>&1
cat gpurun_out/r02_run22_wavefront.jsonl
