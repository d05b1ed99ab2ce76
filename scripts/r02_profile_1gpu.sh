#!/bin/bash
# Final single-GPU run of round 2: whole -m gpu suite, bench lines, ncu launch list + full captures of the hot kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 --durations=10 -rf > gpurun_out/r02_run12_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run12_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run12_bench.json 2> gpurun_out/r02_run12_bench.err
echo "bench rc=$?" >> gpurun_out/r02_run12_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_run12_bench_reference.json 2> gpurun_out/r02_run12_bench_reference.err
timeout 600 python bench.py --size 16384 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r02_run12_bench_n16384.json 2> gpurun_out/r02_run12_bench_n16384.err
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_run12_ensemble.json 2> gpurun_out/r02_run12_ensemble.err
timeout 600 python bench.py --order wavefront --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_run12_bench_wavefront.json 2> gpurun_out/r02_run12_bench_wavefront.err
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_run12_smoke.log 2>&1
export FLUID_BENCH_SETTLE=0
timeout 300 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches_final.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_run12_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_advect_rows|k_rb_stream|k_blur_stream|k_canon_zero" -s 60 -c 14 -o gpurun_out/r02_final_kernels python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-variable-dt > gpurun_out/r02_run12_ncu_full.log 2>&1
tail -6 gpurun_out/r02_run12_tests.log
for f in bench bench_n16384 ensemble bench_wavefront bench_reference; do python - gpurun_out/r02_run12_$f.json <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(sys.argv[1], 'value %.4g'%j['value'], 'ms %.3f'%j['ms_per_step'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
tail -2 gpurun_out/r02_run12_smoke.log
