#!/bin/bash
# where does the FHFMA node test lose? cfg5 at several scene scales, both libraries (development)
for sc in 0.01 0.1 0.25 0.5; do for v in base nofh; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  python bench.py --workload cfg5 --scale $sc --steps 3 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5@$sc $v"
done; done
