#!/bin/bash
# service threshold of the frame programs with the queue kernel (development)
for t in "frame_service_min=8" "frame_service_min=16" "frame_service_min=24" "frame_service_min=32"; do export VSB_TUNE=$t
  for w in cfg2 cfg3 cfg4; do python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-extra 2>/dev/null | python tools/quick.py "$w $t"; done; done
