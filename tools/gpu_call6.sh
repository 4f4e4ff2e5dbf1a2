#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests6.log 2>&1; echo "tests exit $?" >> gpurun_out/r2_gpu_tests6.log; tail -12 gpurun_out/r2_gpu_tests6.log
( time python bench.py ) > gpurun_out/r2_bench_n1b.log 2> gpurun_out/r2_bench_n1b.err; echo "bench exit $?"; tail -4 gpurun_out/r2_bench_n1b.err
python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_n1b.log"):
    if l.startswith("{"):
        d=json.loads(l)
        print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"],1), "frac", round(d["roofline"]["frac"],3))
        for k,v in (d.get("extra") or {}).items():
            if k=="gpu_built_tree":
                for kk,vv in v.items(): print("  gpu tree",kk,round(vv["value"],1),"vs ref tree",vv.get("vs_reference_tree"),vv["builder"])
            else: print("  extra",k,round(v["value"],1),"kernel_ms",round(v["kernel_ms"],3),"frac",round(v["roofline"]["frac"],3), v["builder"])
        print("  builder", d["config"].get("builder"))
PY
