#!/bin/bash
# final multi-GPU lines of the round: bench.py at N = 8 (weak + strong + e2e) and the host D2H ceiling probe
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 \
  > gpurun_out/r2_bench_n${N}_final.log 2> gpurun_out/r2_bench_n${N}_final.err; echo "bench exit $?"
tail -3 gpurun_out/r2_bench_n${N}_final.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/d2h_probe.py \
  > gpurun_out/r2_d2h_probe_n${N}.log 2>&1; echo "probe exit $?"; grep "N=" gpurun_out/r2_d2h_probe_n${N}.log
python - <<PY
import json
for l in open("gpurun_out/r2_bench_n${N}_final.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", d["e2e"])
        for k, v in (d.get("strong") or {}).items():
            print("  strong", k, v["assembly"], round(v["value"], 1), "ms", round(v["ms_per_step"], 3), "eff", round(v["efficiency"], 3), "verified", v["assembled_frame_verified"])
PY
