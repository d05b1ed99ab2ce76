#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631"
timeout 300 $TR scripts/r02_peer_micro.py > gpurun_out/r02_run8_micro.jsonl 2> gpurun_out/r02_run8_micro.err
FLUID_PEER_BLOCKS=16 timeout 300 $TR scripts/r02_peer_micro.py > gpurun_out/r02_run8_micro_b16.jsonl 2> gpurun_out/r02_run8_micro_b16.err
FLUID_PEER_BLOCKS=16 timeout 300 $TR scripts/r02_slab_diag.py 8192 8192 > gpurun_out/r02_run8_diag_b16.jsonl 2> gpurun_out/r02_run8_diag_b16.err
cat gpurun_out/r02_run8_micro.jsonl gpurun_out/r02_run8_micro_b16.jsonl gpurun_out/r02_run8_diag_b16.jsonl; tail -3 gpurun_out/r02_run8_micro.err
