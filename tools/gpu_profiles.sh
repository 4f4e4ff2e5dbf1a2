#!/bin/bash
# ncu evidence for profiles/: launch lists (shares) and one full capture of the dominant kernel per workload.
# Each command has exited 0 WITHOUT ncu in the calls before (gpurun_out/r2_bench_n1c.log).
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-extra"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_cfg2.csv $B > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_cfg5.csv $B --workload cfg5 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 3 --launch-count 1 -o gpurun_out/r2_cfg2_final -f $B > gpurun_out/ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 3 --launch-count 1 -o gpurun_out/r2_cfg5_final -f $B --workload cfg5 > gpurun_out/ncu_c5.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 3 --launch-count 1 -o gpurun_out/r2_cfg5any_final -f $B --workload cfg5 --any > gpurun_out/ncu_c5a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 3 --launch-count 1 -o gpurun_out/r2_cfg3_final -f $B --workload cfg3 > gpurun_out/ncu_c3.log 2>&1
ls -la gpurun_out/*final*.ncu-rep gpurun_out/r2_launches*.csv
