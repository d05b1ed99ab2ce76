#!/bin/bash
# 8 GPUs: weak and strong scaling lines (peer-memory transport), config 4 at 16384 rows per GPU, config 5 on 8 GPUs
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02_run11_gpus.txt 2>&1
run() { # name nproc args...
  name=$1; np=$2; shift 2
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $np "$@" \
     > gpurun_out/r02_run11_$name.json 2> gpurun_out/r02_run11_$name.err
  echo "$name rc=$?" >> gpurun_out/r02_run11_status.txt
}
run weak8192_8gpu 8 --steps 20 --warmup 3
run weak8192_4gpu 4 --steps 20 --warmup 3 --no-e2e
run weak8192_2gpu 2 --steps 20 --warmup 3 --no-e2e
run weak8192_8gpu_nccl 8 --steps 20 --warmup 3 --no-e2e --transport nccl
run weak16384_8gpu 8 --steps 10 --warmup 3 --size 16384 --no-e2e
run strong8192_8gpu 8 --steps 20 --warmup 3 --strong --no-e2e
run strong16384_8gpu 8 --steps 20 --warmup 3 --size 16384 --strong --no-e2e
run strong16384_4gpu 4 --steps 20 --warmup 3 --size 16384 --strong --no-e2e
run strong16384_2gpu 2 --steps 10 --warmup 3 --size 16384 --strong --no-e2e
run ensemble512x512_8gpu 8 --steps 10 --warmup 3 --ensemble 512x512
timeout 600 python -m pytest tests/test_gpu_nccl_slabs.py -q -m gpu -p no:cacheprovider -rf > gpurun_out/r02_run11_nccl_slabs_test.log 2>&1
cat gpurun_out/r02_run11_status.txt
for f in gpurun_out/r02_run11_*.json; do python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(sys.argv[1].split('run11_')[1], 'value %.4g'%j['value'], 'ms %.3f'%j['ms_per_step'], (j.get('slab_check') or '')[:8])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
tail -3 gpurun_out/r02_run11_nccl_slabs_test.log
