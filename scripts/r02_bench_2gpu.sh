#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29651"
for pace in 0 1 2 3; do
  FLUID_PACE_FRAMES=$pace timeout 300 $TR scripts/r02_slab_diag.py 8192 8192 > gpurun_out/r02_run10_diag_pace$pace.jsonl 2> gpurun_out/r02_run10_diag_pace$pace.err
done
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -rf -k "streaming or streamed or ensemble or random_shapes or whole_steps or n8192" > gpurun_out/r02_run10_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run10_tests.log
timeout 600 python scripts/r02_probe3.py ensemble > gpurun_out/r02_run10_ensemble.jsonl 2> gpurun_out/r02_run10_ensemble.err
for pace in 0 1 2 3; do echo pace $pace; grep peer gpurun_out/r02_run10_diag_pace$pace.jsonl; done
tail -5 gpurun_out/r02_run10_tests.log; cut -c1-420 gpurun_out/r02_run10_ensemble.jsonl
