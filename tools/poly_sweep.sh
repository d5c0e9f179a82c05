#!/bin/bash
# A/B of the compacted-pruning polygon compare (poly_compare_f64_queue_kernel) on the GPU box: every
# argument is one set of extra nvcc flags; rebuilds polygon.cu per set, prints kernel ms on c4cap and c4.
set -e
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
OUT=${OUT:-../../gpurun_out/poly_sweep.txt}
: > "$OUT"
for flags in "$@"; do
  rm -f build/polygon.o
  make EXTRA="$flags" -j8 > /dev/null 2>&1
  for wl in c4cap ${C4:+c4}; do
    for k in ${KINDS:-queue legacy}; do
      (cd ../.. && SSP_B200_POLY_COMPARE=$k python bench.py --steps 3 --warmup 3 --workload $wl --no-cpu-baseline 2>/dev/null |
        python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('flags [$flags] $wl $k kernel_ms', round(d['roofline']['kernel_ms_per_launch'],3), 'executed_frac', d['roofline'].get('executed_fraction_of_algorithmic_flops'), 'frac', d['roofline'].get('frac'), 'matches', d['matches_per_step'])") | tee -a "$OUT"
    done
  done
done
rm -f build/polygon.o
make -j8 > /dev/null 2>&1
