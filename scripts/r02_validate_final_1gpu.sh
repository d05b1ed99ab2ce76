#!/bin/bash
# Final single-GPU validation: whole GPU suite, bench lines (main, reference arm, ensemble), smoke
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -rf > gpurun_out/r02_run31_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run31_tests.log
tail -4 gpurun_out/r02_run31_tests.log
FLUID_DEBUG_FEEDBACK=1 timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run31_bench.json 2> gpurun_out/r02_run31_bench.err
echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_run31_bench_reference.json 2> gpurun_out/r02_run31_bench_reference.err
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_run31_ensemble.json 2> gpurun_out/r02_run31_ensemble.err
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_run31_smoke.log 2>&1; echo "smoke rc=$?"
python - <<'PY'
import json
j=json.load(open("gpurun_out/r02_run31_bench.json"))
print("bench", j["ms_per_step"], j["value"], "e2e", j["e2e"]["value"], "variable_dt", j.get("variable_dt"), "clocks", j["clocks"])
j=json.load(open("gpurun_out/r02_run31_ensemble.json")); print("ensemble", j["ms_per_step"], j["value"])
j=json.load(open("gpurun_out/r02_run31_bench_reference.json")); print("reference", j["value"])
PY
grep -c "quick division proof" gpurun_out/r02_run31_bench.err; grep "quick division proof" gpurun_out/r02_run31_bench.err | head -3
tail -2 gpurun_out/r02_run31_smoke.log
