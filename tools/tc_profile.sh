#!/bin/bash
# Cycle accounting of the tcgen05 Hamming kernel (run on the GPU box): rebuilds with -DTC_PROFILE=1,
# block 0 prints where its MMA thread and two epilogue warps spend their cycles, then restores
# the product build.
set -e
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
rm -f build/hamming_tensor.o
make EXTRA=-DTC_PROFILE=1 -j8 > /dev/null
(cd ../.. && python bench.py --steps 1 --warmup 0 --variant ${1:-tensor} --no-cpu-baseline 2>&1 | grep -E "mma thread|epi warp" | head -${2:-16})
rm -f build/hamming_tensor.o
make -j8 > /dev/null
