#!/bin/bash
# ray-list kernel at 9 / 10 CTAs per SM (development)
for v in base l9 l10; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  python bench.py --workload cfg5 --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5 $v" | cut -c1-100
  python bench.py --workload cfg5 --any --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5any $v" | cut -c1-100
done
