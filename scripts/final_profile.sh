#!/bin/bash
# Round-end measurement on one B200: bench line, reference arm, launch list, ncu captures of the hot kernels.
# Everything lands in gpurun_out/final/; summaries are copied into profiles/ by hand afterwards.
set -x
O=gpurun_out/final
mkdir -p $O
python bench.py > $O/bench.json 2> $O/bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$B > $O/b2.json 2> $O/b2.err || exit 1
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file $O/launches.csv $B > $O/ncu_launches.log 2>&1
for spec in "k_rb_stream:60:1:rb_stream" "k_rb_stream:71:1:rb_stream_pow2" "k_rb_stream:70:1:rb_stream_pow2_zero" "k_rb_tile:30:1:rb_tile_redo" "k_blur_stream:4:1:blur_stream" "k_blur_redo:4:1:blur_redo" "k_advect:2:2:advect" "k_divergence:2:1:divergence" "k_subtract_gradient:2:1:gradient" "k_add_source:2:2:add_source" "k_blur_pass:1:1:blur_pass"; do
  IFS=: read k skip cnt name <<< "$spec"
  ncu --set full --import-source on --clock-control none -k regex:$k -s $skip -c $cnt -o $O/$name $B > $O/ncu_$name.log 2>&1
  ncu -i $O/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
  # the reports are ~12 MB each and gpurun brings back 64 MiB at most: keep the two streaming kernels' only
  case $name in rb_stream|blur_stream|rb_stream_pow2) ;; *) rm -f $O/$name.ncu-rep ;; esac
done
python scripts/measure_configs.py > $O/configs.json 2> $O/configs.err
ls -la $O
