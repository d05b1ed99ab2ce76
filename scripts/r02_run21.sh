#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -rf > gpurun_out/r02_run21_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run21_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run21_bench.json 2> gpurun_out/r02_run21_bench.err
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_run21_ensemble.json 2> gpurun_out/r02_run21_ensemble.err
tail -4 gpurun_out/r02_run21_tests.log
python - <<'PY'
import json
for f in ("bench","ensemble"):
    j=json.load(open("gpurun_out/r02_run21_%s.json"%f)); print(f, j["ms_per_step"], j["value"], j.get("classes_ms_per_step") or j.get("breakdown") or "")
PY
