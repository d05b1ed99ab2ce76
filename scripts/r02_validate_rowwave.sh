#!/bin/bash
# Row pipeline in its final form: whole GPU suite, the three reference-order modes by size, timeline of one solve, bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -rf > gpurun_out/r02_run29_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run29_tests.log
tail -4 gpurun_out/r02_run29_tests.log
timeout 900 python scripts/r02_probe3.py wavefront_rows > gpurun_out/r02_run29_wavefront_modes.jsonl 2>&1
cut -c1-120 gpurun_out/r02_run29_wavefront_modes.jsonl
timeout 300 python scripts/r02_rowwave_stamps_summary.py 1024 > gpurun_out/r02_run29_timeline_n1024.txt 2>&1
timeout 300 python scripts/r02_rowwave_stamps_summary.py 8192 > gpurun_out/r02_run29_timeline_n8192.txt 2>&1
tail -6 gpurun_out/r02_run29_timeline_n1024.txt gpurun_out/r02_run29_timeline_n8192.txt
timeout 300 python scripts/r02_rowwave_micro.py 8192 > gpurun_out/r02_run29_micro.txt 2>&1
timeout 600 python bench.py --size 1024 --order wavefront --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_run29_bench_wavefront_n1024.json 2> gpurun_out/r02_run29_bench_wavefront_n1024.err
cut -c1-250 gpurun_out/r02_run29_bench_wavefront_n1024.json
