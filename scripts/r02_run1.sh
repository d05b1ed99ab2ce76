#!/bin/bash
# GPU run 1 of round 2: full -m gpu suite, bench line, advect probe, ncu capture of the row-window advect kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02_run1_gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --durations=15 > gpurun_out/r02_run1_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run1_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run1_bench.json 2> gpurun_out/r02_run1_bench.err
echo "bench rc=$?" >> gpurun_out/r02_run1_bench.err
timeout 600 python scripts/r02_advect_probe.py 8192 10 > gpurun_out/r02_run1_advect_probe.jsonl 2> gpurun_out/r02_run1_advect_probe.err
rc=$?
echo "probe rc=$rc" >> gpurun_out/r02_run1_advect_probe.err
if [ $rc -eq 0 ]; then
  timeout 600 python scripts/r02_advect_probe.py 8192 1 --quick > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_advect_rows -c 2 -o gpurun_out/r02_advect_rows python scripts/r02_advect_probe.py 8192 1 --quick > gpurun_out/r02_run1_ncu.log 2>&1
fi
tail -5 gpurun_out/r02_run1_tests.log
cat gpurun_out/r02_run1_bench.json | head -c 1500
cat gpurun_out/r02_run1_advect_probe.jsonl
