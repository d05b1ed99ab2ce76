#!/bin/bash
# 2 GPUs: MF tests, multi-rank slab tests (NCCL + peer transport), 2-GPU bench lines with both transports
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02_run5_gpus.txt 2>&1
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -rf -k "multifluid or nccl or multi or slab or advect" > gpurun_out/r02_run5_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run5_tests.log
for tr in peer nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 3 --transport $tr > gpurun_out/r02_run5_bench_2gpu_$tr.json 2> gpurun_out/r02_run5_bench_2gpu_$tr.err
  echo "bench $tr rc=$?" >> gpurun_out/r02_run5_bench_2gpu_$tr.err
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 3 --strong --no-e2e > gpurun_out/r02_run5_bench_2gpu_strong.json 2> gpurun_out/r02_run5_bench_2gpu_strong.err
tail -8 gpurun_out/r02_run5_tests.log
for f in gpurun_out/r02_run5_bench_2gpu_*.json; do python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(sys.argv[1], j['value'], j['ms_per_step'], j.get('slab_check'), j['roofline']['classes_ms_per_step'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
tail -5 gpurun_out/r02_run5_bench_2gpu_peer.err
