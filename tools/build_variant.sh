#!/bin/bash
# tools/build_variant.sh NAME [-DVSB_...=...]...   -> visionaray_b200/lib/variants/libvsnray_b200_NAME.so
# Tuning variants of the CUDA library (same sources, different compile-time constants); load one
# with VSB_CUDA_LIB=<path>.  Development only.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p visionaray_b200/lib/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off -shared \
     -Iinclude "$@" visionaray_b200/csrc/capi.cu -o visionaray_b200/lib/variants/libvsnray_b200_$name.so
echo built $name
