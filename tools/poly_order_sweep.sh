#!/bin/bash
# Which property group should kernel 3's exact pruning examine first?  (run on the GPU box)
set -e
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
for o in "0 1" "1 0" "2 1" "1 2" "2 0" "0 2"; do
  set -- $o
  rm -f build/polygon.o
  make EXTRA="-DPOLY_G0=$1 -DPOLY_G1=$2" -j8 > /dev/null 2>&1
  for wl in c4 c4cap; do
    (cd ../.. && python bench.py --workload $wl --steps 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('order $1,$2 $wl kernel_ms', round(d['roofline']['kernel_ms_per_launch'],3), 'matches', d['matches_per_step'])")
  done
done
rm -f build/polygon.o
make -j8 > /dev/null 2>&1
