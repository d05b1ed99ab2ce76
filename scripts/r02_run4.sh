#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 --durations=8 -rf > gpurun_out/r02_run4_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run4_tests.log
timeout 600 python scripts/r02_advect_probe.py 8192 10 > gpurun_out/r02_run4_advect_probe.jsonl 2> gpurun_out/r02_run4_advect_probe.err
timeout 900 python scripts/r02_probe3.py wavefront > gpurun_out/r02_run4_probe.jsonl 2> gpurun_out/r02_run4_probe.err
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run4_bench.json 2> gpurun_out/r02_run4_bench.err
tail -12 gpurun_out/r02_run4_tests.log
cat gpurun_out/r02_run4_advect_probe.jsonl gpurun_out/r02_run4_probe.jsonl
