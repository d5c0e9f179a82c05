#!/bin/bash
# A/B sweeps of the tcgen05 Hamming kernel on the GPU box: every argument is one set of extra
# nvcc flags (e.g. "-DTC_FULL_MIN=5" or "-DTC_EXPERIMENT=2"; "" = product build).  Rebuilds
# libssp_b200.so per set, prints the dominant kernel's ms per launch on C3 (and C5 with C5=1),
# then restores the product build.  TC_EXPERIMENT > 0 gives wrong results by design.
set -e
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
OUT=${OUT:-../../gpurun_out/tc_sweep.txt}
: > "$OUT"
for flags in "$@"; do
  rm -f build/hamming_tensor.o
  make EXTRA="$flags" -j8 > /dev/null 2>&1
  for wl in c3 ${C5:+c5}; do
    (cd ../.. && python bench.py --steps 3 --warmup 2 --workload $wl --no-cpu-baseline --no-sub 2>/dev/null |
      python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('flags [$flags] $wl kernel_ms', round(d['roofline']['kernel_ms_per_launch'],2), 'sm_mhz', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'matches', d['matches_per_step'])") | tee -a "$OUT"
  done
done
rm -f build/hamming_tensor.o
make -j8 > /dev/null 2>&1
