#!/bin/bash
# experimental builds of the scaled-sequence streaming kernels on the 512 x N=512 ensemble
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for v in "" v1 v2 v4 v5; do
  if [ -n "$v" ]; then export FLUID_ALT_LIB=fluid-simulation_b200/build_alt/libfluid_$v.so; fi
  timeout 300 python scripts/r02_ens_list.py 2 12 5 2>&1 | tail -1
done | tee gpurun_out/r02_run18_variants.log
