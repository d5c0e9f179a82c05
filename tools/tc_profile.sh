#!/bin/bash
# Cycle accounting of the tcgen05 Hamming kernel (run on the GPU box): rebuilds with -DTC_PROFILE=1,
# block 0 prints where its MMA thread and two epilogue warps spend their cycles and, from an
# observer warp, the mean time per train tile by position inside the work item; then restores
# the product build.   usage: tc_profile.sh [workload=c3] [lines=40] [extra nvcc flags]
set -e
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
rm -f build/hamming_tensor.o
make EXTRA="-DTC_PROFILE=1 $3" -j8 > /dev/null 2>&1
(cd ../.. && python bench.py --steps 1 --warmup 0 --workload ${1:-c3} --no-cpu-baseline 2>&1 | grep -E "mma thread|epi warp|stage time" | head -${2:-40})
rm -f build/hamming_tensor.o
make -j8 > /dev/null 2>&1
