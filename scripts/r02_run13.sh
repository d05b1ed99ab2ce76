#!/bin/bash
# Scaled division sequence (exact on the subnormal fringe): parity tests of the streaming kernels, ensemble probe and bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -x -k "stream or ensemble or blur or division" -rf > gpurun_out/r02_run13_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run13_tests.log
FLUID_DEBUG_FEEDBACK=1 timeout 600 python scripts/r02_probe3.py ensemble_modes > gpurun_out/r02_run13_ensemble_modes.jsonl 2> gpurun_out/r02_run13_ensemble_modes.err
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_run13_ensemble.json 2> gpurun_out/r02_run13_ensemble.err
timeout 600 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_run13_bench.json 2> gpurun_out/r02_run13_bench.err
tail -5 gpurun_out/r02_run13_tests.log
cat gpurun_out/r02_run13_ensemble_modes.jsonl
tail -3 gpurun_out/r02_run13_ensemble_modes.err | cut -c1-300
cut -c1-400 gpurun_out/r02_run13_ensemble.json
cut -c1-300 gpurun_out/r02_run13_bench.json
