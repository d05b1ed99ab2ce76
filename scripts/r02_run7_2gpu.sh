#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 scripts/r02_slab_diag.py 8192 8192 > gpurun_out/r02_run7_diag.jsonl 2> gpurun_out/r02_run7_diag.err
echo "rc=$?" >> gpurun_out/r02_run7_diag.err
timeout 600 python scripts/r02_probe3.py ensemble > gpurun_out/r02_run7_ensemble.jsonl 2> gpurun_out/r02_run7_ensemble.err
cat gpurun_out/r02_run7_diag.jsonl gpurun_out/r02_run7_ensemble.jsonl; tail -3 gpurun_out/r02_run7_diag.err
