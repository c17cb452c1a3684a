#!/bin/bash
# Round-1 evidence capture on one B200 (run under gpurun): launch list, then one `--set full` capture per hot kernel,
# all at the bench's full size (10 000 targets).  Numbers printed under ncu are never bench values.
set -u
B="python bench.py --steps 1 --warmup 1 --no-cpu"
$B > gpurun_out/bench_pre_ncu.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_10000targets.csv $B > gpurun_out/ncu_launch.log 2>&1
for k in k_merge_best2 k_build_graph2 k_normalize_warp; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o gpurun_out/full10k_$k $B > gpurun_out/ncu_full10k_$k.log 2>&1
done
ls -la gpurun_out
