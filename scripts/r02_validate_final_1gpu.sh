#!/bin/bash
# Final single-GPU validation: whole GPU suite, bench lines (main, reference arm, ensemble, reference order at N=1024), smoke
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -rf > gpurun_out/r02_final_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_final_tests.log
tail -4 gpurun_out/r02_final_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err
echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_final_bench_reference.json 2> gpurun_out/r02_final_bench_reference.err
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_final_ensemble.json 2> gpurun_out/r02_final_ensemble.err
timeout 600 python bench.py --size 1024 --order wavefront --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_final_bench_wavefront_n1024.json 2> gpurun_out/r02_final_bench_wavefront_n1024.err
timeout 600 python bench.py --order wavefront --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_final_bench_wavefront.json 2> gpurun_out/r02_final_bench_wavefront.err
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_final_smoke.log 2>&1; echo "smoke rc=$?"
python - <<'PY'
import json
j=json.load(open("gpurun_out/r02_final_bench.json"))
print("bench", j["ms_per_step"], j["value"], "e2e", j["e2e"]["value"], "variable_dt", j["variable_dt"]["ms_per_step"], "clocks", j["clocks"])
for f in ("ensemble","bench_wavefront_n1024","bench_wavefront","bench_reference"):
    j=json.load(open("gpurun_out/r02_final_%s.json"%f)); print(f, j.get("ms_per_step"), j["value"])
PY
tail -2 gpurun_out/r02_final_smoke.log
