#!/bin/bash
# prefetch of a group's records at append time (development)
for v in base pf1 pf2; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  for w in cfg3 cfg4; do python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-extra 2>/dev/null | python tools/quick.py "$w $v" | cut -c1-100; done
  python bench.py --workload cfg5 --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5 $v" | cut -c1-100
  python bench.py --workload cfg5 --any --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5any $v" | cut -c1-100
done
