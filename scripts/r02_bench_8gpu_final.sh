#!/bin/bash
# 8 GPUs, final build: the driver's scaling sequence (1, 2, 4, 8 with default arguments), config 5 on 8 GPUs, strong 8192, the multi-rank test
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run() { # name nproc args...
  name=$1; np=$2; shift 2
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $np "$@" \
     > gpurun_out/r02_run30_$name.json 2> gpurun_out/r02_run30_$name.err
  echo "$name rc=$?" >> gpurun_out/r02_run30_status.txt
}
timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02_run30_weak8192_1gpu.json 2> gpurun_out/r02_run30_weak8192_1gpu.err
run weak8192_2gpu 2 --steps 20 --warmup 3
run weak8192_4gpu 4 --steps 20 --warmup 3
run weak8192_8gpu 8 --steps 20 --warmup 3
run ensemble512x512_8gpu 8 --steps 10 --warmup 3 --ensemble 512x512
run strong8192_8gpu 8 --steps 20 --warmup 3 --strong --no-e2e
timeout 600 python -m pytest tests/test_gpu_nccl_slabs.py -q -m gpu -p no:cacheprovider -rf > gpurun_out/r02_run30_nccl_slabs_test.log 2>&1
cat gpurun_out/r02_run30_status.txt
for f in gpurun_out/r02_run30_*.json; do python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(sys.argv[1].split('run30_')[1], 'value %.4g'%j['value'], 'ms %.3f'%j['ms_per_step'], (j.get('slab_check') or '')[:8])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
tail -3 gpurun_out/r02_run30_nccl_slabs_test.log
