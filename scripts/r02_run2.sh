#!/bin/bash
# GPU run 2 of round 2: full -m gpu suite with durations, bench line, advect probe, memcheck on a subset
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 --durations=25 -rf > gpurun_out/r02_run2_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run2_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_run2_bench.json 2> gpurun_out/r02_run2_bench.err
echo "bench rc=$?" >> gpurun_out/r02_run2_bench.err
timeout 600 python scripts/r02_advect_probe.py 8192 10 > gpurun_out/r02_run2_advect_probe.jsonl 2> gpurun_out/r02_run2_advect_probe.err
echo "probe rc=$?" >> gpurun_out/r02_run2_advect_probe.err
timeout 600 python bench.py --ensemble 512x512 --steps 10 --warmup 3 > gpurun_out/r02_run2_ensemble.json 2> gpurun_out/r02_run2_ensemble.err
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests -q -m gpu -p no:cacheprovider -x \
  -k "streaming_interior_kernel_bitwise or streamed_blur_bitwise or test_small_grid or advect_row_window_kernel_bitwise or sparse_add or bordered_primitives or test_vel_and_dens_step or large_slabs_use" \
  > gpurun_out/r02_run2_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/r02_run2_memcheck.log
tail -40 gpurun_out/r02_run2_tests.log
cat gpurun_out/r02_run2_advect_probe.jsonl
tail -3 gpurun_out/r02_run2_memcheck.log
