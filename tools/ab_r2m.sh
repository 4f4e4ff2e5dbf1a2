#!/bin/bash
# instruction counts and stall picture of the ray-list kernel, FHFMA vs f32 node test (development)
M=smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp16.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum
for v in base fh; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  ncu --metrics $M --clock-control none -k regex:trace_kernel --launch-skip 3 --launch-count 1 --csv --log-file gpurun_out/_m_$v.csv python bench.py --workload cfg5 --scale 0.5 --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/_m_$v.csv")) if len(r)>5]
h=rows[0]
for r in rows[1:]:
    print("$v", r[h.index("Metric Name")], r[h.index("Metric Value")])
PY
done
