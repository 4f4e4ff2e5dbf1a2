#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_reference_cuda.py tests/test_ranges_and_streams.py -m gpu -x -q 2>&1 | tail -4
( time python bench.py ) > gpurun_out/r2_bench_n1c.log 2> gpurun_out/r2_bench_n1c.err; echo "bench exit $?"; tail -4 gpurun_out/r2_bench_n1c.err
python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_n1c.log"):
    if l.startswith("{"):
        d=json.loads(l)
        print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"],1))
        print("naive_gpu", d["extra"].get("naive_gpu"))
PY
