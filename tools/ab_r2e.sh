#!/bin/bash
V=$PWD/visionaray_b200/lib/variants/libvsnray_b200_gring.so
VSB_CUDA_LIB=$V python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_surface.py -m gpu -x -q 2>&1 | tail -3
for v in base gring base gring; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$V; fi
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra 2>&1 | python tools/quick.py "cfg2 $v"
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra --workload cfg4 --builder gpu 2>&1 | python tools/quick.py "cfg4 $v"
  python bench.py --steps 5 --warmup 3 --no-cpu --no-extra --workload cfg3 --builder gpu 2>&1 | python tools/quick.py "cfg3 $v"
  python bench.py --workload cfg5 --scale 0.25 --batch 8388608 --steps 5 --warmup 3 --no-cpu 2>&1 | python tools/quick.py "cfg5/4 $v"
  python bench.py --workload cfg5 --scale 0.25 --batch 8388608 --steps 5 --warmup 3 --no-cpu --any 2>&1 | python tools/quick.py "cfg5/4 any $v"
done
