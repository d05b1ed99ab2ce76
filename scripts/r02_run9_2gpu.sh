#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641"
timeout 300 $TR scripts/r02_slab_series.py > gpurun_out/r02_run9_series.jsonl 2> gpurun_out/r02_run9_series.err
FLUID_BENCH_NO_SAMPLER=1 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --transport nccl --no-e2e > gpurun_out/r02_run9_bench_nosampler.json 2> gpurun_out/r02_run9_bench_nosampler.err
FLUID_BENCH_SAMPLE_MS=200 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --transport nccl --no-e2e > gpurun_out/r02_run9_bench_s200.json 2> gpurun_out/r02_run9_bench_s200.err
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --transport nccl --no-e2e > gpurun_out/r02_run9_bench_s25.json 2> gpurun_out/r02_run9_bench_s25.err
timeout 300 $TR bench.py --gpus 2 --steps 60 --warmup 10 --transport nccl --no-e2e > gpurun_out/r02_run9_bench_s25_k60.json 2> gpurun_out/r02_run9_bench_s25_k60.err
cat gpurun_out/r02_run9_series.jsonl
for f in gpurun_out/r02_run9_bench_*.json; do python -c "
import json,sys
j=json.load(open('$f')); print('$f', j['ms_per_step'], j['wall_ms_per_step'], j['clocks'])"; done
