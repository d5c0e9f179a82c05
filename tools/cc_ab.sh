#!/bin/bash
# debug: sparse vs dense cross-check, for each set of nvcc flags given
cd "$(dirname "$0")/.."
mkdir -p /tmp/cc
for flags in "$@"; do
  (cd opencv-starry-sky-panorama-stitcher_b200/csrc && rm -f build/hamming_tensor.o && make EXTRA="$flags" -j8 >/dev/null 2>&1)
  echo "== flags [$flags]"
  SSP_B200_CROSSCHECK_DENSE=1 timeout 150 python tools/cc_debug.py ${NIMG:-21} 50000 /tmp/cc/cc_dense.npz 2>&1 | tail -3
  SSP_B200_CROSSCHECK_DENSE=0 timeout 150 python tools/cc_debug.py ${NIMG:-21} 50000 /tmp/cc/cc_sparse.npz 2>&1 | tail -3
  python tools/cc_diff.py
done
(cd opencv-starry-sky-panorama-stitcher_b200/csrc && rm -f build/hamming_tensor.o && make -j8 >/dev/null 2>&1)
