#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/r2_gpu_tests4.log 2>&1; echo "tests exit $?" >> gpurun_out/r2_gpu_tests4.log
tail -12 gpurun_out/r2_gpu_tests4.log
( time python bench.py ) > gpurun_out/r2_bench_n1.log 2> gpurun_out/r2_bench_n1.err; echo "bench exit $?"; tail -3 gpurun_out/r2_bench_n1.err
( time python bench.py --impl reference --steps 10 --warmup 3 ) > gpurun_out/r2_bench_ref.log 2> gpurun_out/r2_bench_ref.err; echo "ref exit $?"; tail -3 gpurun_out/r2_bench_ref.err
python - <<'PY'
import json
for f in ("gpurun_out/r2_bench_n1.log","gpurun_out/r2_bench_ref.log"):
    for l in open(f):
        if l.startswith("{"):
            d=json.loads(l)
            print(f, "value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"],1), "frame", d["config"].get("frame"))
            for k,v in (d.get("extra") or {}).items(): print("  extra",k,round(v["value"],1),"kernel_ms",round(v["kernel_ms"],3),"frac",round(v["roofline"]["frac"],3))
            print("  builder", d["config"].get("builder"))
PY
