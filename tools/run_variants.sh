# usage: bash tools/run_variants.sh "<VSB_TUNE string or ->" variant...   (variant "base" = the in-tree library)
tune=$1; shift
for v in "$@"; do
  if [ "$v" = base ]; then unset VSB_CUDA_LIB; else export VSB_CUDA_LIB=$PWD/visionaray_b200/lib/variants/libvsnray_b200_$v.so; fi
  if [ "$tune" != "-" ]; then export VSB_TUNE=$tune; fi
  python bench.py --workload cfg5 --scale 0.25 --batch 8388608 --steps 3 --warmup 3 --no-cpu 2>&1 | python tools/quick.py "cfg5 $v $tune"
  python bench.py --steps 3 --warmup 3 --no-cpu 2>&1 | python tools/quick.py "cfg2 $v $tune"
done
