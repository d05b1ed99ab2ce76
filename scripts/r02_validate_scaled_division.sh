#!/bin/bash
# Scaled division sequence, final form: whole GPU suite, ensemble on automatic, modes at N=8192 and on ensembles, bench lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 --durations=8 -rf > gpurun_out/r02_run20_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run20_tests.log
FLUID_DEBUG_FEEDBACK=1 timeout 300 python scripts/r02_ens_auto.py > gpurun_out/r02_run20_ens_auto.log 2>&1
timeout 600 python scripts/r02_modes.py 8192 1 10 > gpurun_out/r02_run20_modes.jsonl 2>&1
timeout 600 python scripts/r02_modes.py 512 512 10 >> gpurun_out/r02_run20_modes.jsonl 2>&1
timeout 600 python scripts/r02_modes.py 1024 128 10 >> gpurun_out/r02_run20_modes.jsonl 2>&1
timeout 600 python scripts/r02_modes.py 2048 16 10 >> gpurun_out/r02_run20_modes.jsonl 2>&1
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_run20_ensemble.json 2> gpurun_out/r02_run20_ensemble.err
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run20_bench.json 2> gpurun_out/r02_run20_bench.err
tail -12 gpurun_out/r02_run20_tests.log
cat gpurun_out/r02_run20_ens_auto.log
cut -c1-330 gpurun_out/r02_run20_modes.jsonl
cut -c1-300 gpurun_out/r02_run20_ensemble.json
cut -c1-300 gpurun_out/r02_run20_bench.json
