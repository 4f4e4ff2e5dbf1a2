#!/bin/bash
# FHFMA node test against the f32 one, interleaved (development)
for rep in 1 2; do for v in base nofh; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  bash tools/ab_r2j.sh $v
done; done
