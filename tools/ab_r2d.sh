#!/bin/bash
# A/B after passing the scene record by value: base (6 CTAs / SM) vs f5 (frame kernels at 5 CTAs / SM, 96 registers)
for v in base f5 base f5; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra 2>&1 | python tools/quick.py "cfg2 $v"
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra --workload cfg4 --builder gpu 2>&1 | python tools/quick.py "cfg4 $v"
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra --workload cfg3 --builder gpu 2>&1 | python tools/quick.py "cfg3 $v"
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra --workload cfg1 2>&1 | python tools/quick.py "cfg1 $v"
done
unset VSB_CUDA_LIB
python bench.py --workload cfg5 --scale 0.25 --batch 8388608 --steps 5 --warmup 3 --no-cpu 2>&1 | python tools/quick.py "cfg5/4 base"
