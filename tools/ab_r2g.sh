#!/bin/bash
# occupancy variants and drain thresholds with the round-2 node test (development)
mkdir -p gpurun_out
run() { tag=$1; shift; for w in "$@"; do
  if [ "$w" = cfg5 ]; then python bench.py --workload cfg5 --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5 $tag";
  elif [ "$w" = cfg5any ]; then python bench.py --workload cfg5 --any --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5any $tag";
  else python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-extra 2>/dev/null | python tools/quick.py "$w $tag"; fi; done; }
for v in f5 f7; do export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; run $v cfg2 cfg3 cfg4; done
export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_l6.so; run l6 cfg5 cfg5any
unset VSB_CUDA_LIB
for t in "drain_min=16" "drain_min=32" "pend_min=4" "pend_min=16" "drain_min=28,pend_min=12"; do export VSB_TUNE=$t; run "$t" cfg2 cfg4 cfg5; done
