#!/bin/bash
# all workloads once (development)
for w in cfg1 cfg2 cfg3 cfg4; do python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-extra 2>/dev/null | python tools/quick.py "$w $1"; done
python bench.py --workload cfg5 --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5 $1"
python bench.py --workload cfg5 --any --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5any $1"
