#!/bin/bash
# service threshold of the ray-list program with the group ring (development)
for w in cfg3 cfg4; do python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-extra 2>/dev/null | python tools/quick.py "$w default"; done
for t in "query_service_min=2" "query_service_min=4" "query_service_min=8" "query_service_min=16"; do export VSB_TUNE=$t
  python bench.py --workload cfg5 --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5 $t"
  python bench.py --workload cfg5 --any --steps 5 --warmup 3 --no-cpu 2>/dev/null | python tools/quick.py "cfg5any $t"; done
