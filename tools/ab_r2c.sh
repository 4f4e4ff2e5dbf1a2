#!/bin/bash
# round 2, call 3 (after the machine was replaced): full GPU tier, then base vs hcvt, then ncu captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/r2_gpu_tests3.log 2>&1; echo "tests exit $?" >> gpurun_out/r2_gpu_tests3.log
tail -14 gpurun_out/r2_gpu_tests3.log
{
bash tools/run_variants.sh - base hcvt base hcvt
} > gpurun_out/r2_ab3.log 2>&1
cat gpurun_out/r2_ab3.log
ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r2_cfg2_q2 -f python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:trace_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r2_cfg5_q2 -f python bench.py --workload cfg5 --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_c5.log 2>&1
ls -la gpurun_out/*.ncu-rep
