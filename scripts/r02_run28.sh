#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
for v in "" b c; do
  if [ -n "$v" ]; then export FLUID_ALT_LIB=fluid-simulation_b200/build_alt/libfluid_$v.so; fi
  echo "== variant ${v:-main}"
  timeout 300 python scripts/r02_rowwave_micro.py 8192 2>&1 | head -3
  for n in 1024 2048 8192; do timeout 120 python scripts/r02_rowwave_one.py $n 3 3; done 2>&1
done
