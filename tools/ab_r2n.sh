#!/bin/bash
# e2e of the headline frame against the number of bands (development)
for t in "frame_bands=8" "frame_bands=12" "frame_bands=16" "frame_bands=16,band_min_mib=8" "frame_bands=24,band_min_mib=8" "frame_bands=32,band_min_mib=4"; do export VSB_TUNE=$t
  python bench.py --workload cfg2 --steps 5 --warmup 3 --no-cpu --no-extra 2>/dev/null | python tools/quick.py "cfg2 $t" | cut -c1-150; done
