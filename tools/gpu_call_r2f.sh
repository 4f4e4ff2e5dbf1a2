#!/bin/bash
# full GPU validation of the round: all -m gpu tests, smoke, default bench + reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests_final.log 2>&1; echo "tests exit $?" >> gpurun_out/r2_gpu_tests_final.log; tail -6 gpurun_out/r2_gpu_tests_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
( time python bench.py ) > gpurun_out/r2_bench_final.log 2> gpurun_out/r2_bench_final.err; echo "bench exit $?"; tail -4 gpurun_out/r2_bench_final.err
( time python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/r2_bench_ref_final.log 2> gpurun_out/r2_bench_ref_final.err; echo "ref exit $?"; tail -4 gpurun_out/r2_bench_ref_final.err; tail -c 600 gpurun_out/r2_bench_ref_final.log
python - <<'PY'
import json
for l in open("gpurun_out/r2_bench_final.log"):
    if l.startswith("{"):
        d=json.loads(l)
        print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"],1), "frac", round(d["roofline"]["frac"],3))
        print("  builder", d["config"].get("builder"))
        for k,v in (d.get("extra") or {}).items():
            if k=="ploc_tree":
                for kk,vv in v.items(): print("  ploc tree",kk,round(vv["value"],1),"vs headline tree",vv.get("vs_headline_tree"),vv["builder"])
            elif k=="naive_gpu": print("  naive_gpu", v)
            else: print("  extra",k,round(v["value"],1),"kernel_ms",round(v["kernel_ms"],3),"frac",round(v["roofline"]["frac"],3), v["builder"])
PY
