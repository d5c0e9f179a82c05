#!/bin/bash
# Bottleneck probes of the tcgen05 Hamming kernel (run on the GPU box: nvcc is in the image).
# Rebuilds libssp_b200.so with -DTC_EXPERIMENT=n (results are WRONG for n > 0; only the time of the
# dominant kernel is of interest), prints kernel ms per launch, then restores the product build.
#   1: accumulators are never read      -> TMA feed + MMA issue rate alone
#   2, 3: pass 1 only (packed filter), no exact updates
#   4: like 1 with 0/1 instead of +-1 operands -> operand-toggling share of the power cap
set -e
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
OUT=${1:-../../gpurun_out/tc_bottleneck.txt}
: > "$OUT"
for n in ${EXPERIMENTS:-1 2 3 0}; do
  rm -f build/hamming_tensor.o
  make EXTRA=-DTC_EXPERIMENT=$n -j8 > /dev/null
  for v in ${VARIANTS:-tensor}; do
    (cd ../.. && python bench.py --steps 3 --warmup 2 --variant $v --no-cpu-baseline --no-sub 2>/dev/null |
      python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('experiment $n variant $v kernel_ms', round(d['roofline']['kernel_ms_per_launch'],2), 'sm_mhz', d['clocks']['sm_mhz'], d['clocks']['reasons'])") | tee -a "$OUT"
  done
done
rm -f build/hamming_tensor.o
make -j8 > /dev/null
