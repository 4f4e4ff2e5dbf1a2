#!/bin/bash
# round 2, call 1: parity with the queue kernel as default, then A/B of trace_kernel vs trace_kernel_q
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r2_gpu_tests.log
tail -3 gpurun_out/r2_gpu_tests.log
{
for tune in trace_queue=0 trace_queue=1 trace_queue=1,drain_min=24 trace_queue=1,pend_min=4 trace_queue=1,pend_min=16 trace_queue=1,pend_min=32 trace_queue=1,query_service_min=8 trace_queue=1,frame_service_min=24; do
  bash tools/run_variants.sh $tune base
done
bash tools/run_variants.sh trace_queue=1 q8
bash tools/run_variants.sh trace_queue=1,pend_min=16 q8
} > gpurun_out/r2_ab1.log 2>&1
cat gpurun_out/r2_ab1.log
