#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests5.log 2>&1; echo "tests exit $?" >> gpurun_out/r2_gpu_tests5.log; tail -3 gpurun_out/r2_gpu_tests5.log
{
for t in build_cost_prim_x100=100 build_cost_prim_x100=70 build_cost_prim_x100=50 build_cost_prim_x100=30 build_cost_node_x100=200 build_cost_node_x100=300,build_cost_prim_x100=70 build_radius=16 build_radius=16,build_cost_prim_x100=50; do
  export VSB_TUNE=$t
  python bench.py --no-cpu --no-extra --steps 5 2>&1 | python tools/quick.py "cfg2 gpu $t"
  python bench.py --workload cfg5 --scale 0.25 --batch 8388608 --no-cpu --steps 5 2>&1 | python tools/quick.py "cfg5 gpu $t"
done
unset VSB_TUNE
python bench.py --no-cpu --no-extra --steps 5 --builder reference 2>&1 | python tools/quick.py "cfg2 reference tree"
python bench.py --workload cfg5 --scale 0.25 --batch 8388608 --no-cpu --steps 5 --builder reference 2>&1 | python tools/quick.py "cfg5 reference tree"
} > gpurun_out/r2_builder_sweep.log 2>&1
cat gpurun_out/r2_builder_sweep.log
