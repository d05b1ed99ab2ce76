#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --timeout=400 -x -k "stream or ensemble or blur or division" -rf > gpurun_out/r02_run19_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_run19_tests.log
timeout 600 python scripts/r02_probe3.py ensemble_modes > gpurun_out/r02_run19_ensemble_modes.jsonl 2> gpurun_out/r02_run19_ensemble_modes.err
timeout 300 python scripts/r02_ens_list.py 2 12 > gpurun_out/r02_run19_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_run19_ens_list_mode2.csv python scripts/r02_ens_list.py 2 12 > gpurun_out/r02_run19_ncu.log 2>&1
tail -3 gpurun_out/r02_run19_tests.log
cut -c1-420 gpurun_out/r02_run19_ensemble_modes.jsonl
tail -1 gpurun_out/r02_run19_plain.log
