#!/bin/bash
# ncu evidence for profiles/: launch lists (shares) and one full capture of the dominant kernel per workload, summarised ON
# the box (gpurun copies at most 64 MiB back).  Each command has exited 0 WITHOUT ncu before (gpurun_out/r2_ab7.log, r2_bench_lut.log).
mkdir -p gpurun_out /tmp/reps
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-extra"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2c_launches_cfg2.csv $B > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2c_launches_cfg5.csv $B --workload cfg5 > /dev/null 2>&1
cap() { name=$1; shift; ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 3 --launch-count 1 -o /tmp/reps/$name -f $B "$@" > gpurun_out/ncu_$name.log 2>&1
  python tools/ncu_summary.py raw /tmp/reps/$name.ncu-rep > gpurun_out/${name}_raw.txt 2>&1
  python tools/ncu_groups.py /tmp/reps/$name.ncu-rep > gpurun_out/${name}_regions.txt 2>&1
  python tools/ncu_opcodes.py /tmp/reps/$name.ncu-rep 28 > gpurun_out/${name}_opcodes.txt 2>&1
  for r in "" "node_test" "queue: append" "queue: drain" "node phase" "kernel loop" "isect_triangle" "shade"; do python tools/ncu_region_opcodes.py /tmp/reps/$name.ncu-rep "$r" 2>&1 | head -16; done > gpurun_out/${name}_region_opcodes.txt
  ncu -i /tmp/reps/$name.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
r=list(csv.reader(sys.stdin)); h,u,row=r[0],r[1],r[2]
for i,n in enumerate(h):
    if n.startswith('sm__inst_executed_pipe') and n.endswith('avg.pct_of_peak_sustained_active'): print(n, row[i])
" > gpurun_out/${name}_pipes.txt; }
cap r2c_cfg2
cap r2c_cfg5 --workload cfg5
cap r2c_cfg5any --workload cfg5 --any
cap r2c_cfg3 --workload cfg3
cap r2c_cfg4 --workload cfg4
cp /tmp/reps/r2c_cfg2.ncu-rep gpurun_out/
ls -la gpurun_out/
