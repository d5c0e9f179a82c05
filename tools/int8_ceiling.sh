#!/bin/bash
# Measured int8 ceiling of kernel 1's tensor path (run on the GPU box: nvcc is in the image).
# Builds libssp_b200.so with -DTC_EXPERIMENT=1 -- the product kernel with the epilogue compiled out: TMA
# feed + tcgen05.mma issue + accumulator hand-back, nothing else; results are wrong by design -- times
# the C3 launch with bench.py, writes profiles/int8_ceiling_rNN.json with the clock record, restores the
# product build.  bench.py quotes roofline.frac against this number.
set -e
ROUND=${ROUND:-r02}
cd "$(dirname "$0")/../opencv-starry-sky-panorama-stitcher_b200/csrc"
rm -f build/hamming_tensor.o
make EXTRA=-DTC_EXPERIMENT=1 -j8 > /dev/null
(cd ../.. && python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sub 2>/dev/null | python -c "
import json, sys
d = json.loads(sys.stdin.read())
r = d['roofline']
ms = r['kernel_ms_per_launch']
tops = r['comparisons_per_launch'] * 512 / (ms / 1e3) / 1e12
out = {'what': 'MMA-only probe: hamming_tcgen05_kernel<2> built with -DTC_EXPERIMENT=1 (epilogue compiled out), C3 launch '
               '(210 pairs x 50k x 50k), mean of 10 launches inside bench.py, L2 flushed between launches',
       'kernel_ms': ms, 'tops': tops, 'comparisons_per_launch': r['comparisons_per_launch'],
       'sm_mhz': d['clocks']['sm_mhz'], 'sm_max_mhz': d['clocks']['sm_max_mhz'], 'reasons': d['clocks']['reasons'],
       'pipe_peak_at_max_clock': r['pipe_peak_at_max_clock']}
json.dump(out, open('profiles/int8_ceiling_${ROUND}.json', 'w'), indent=1)
print(json.dumps(out))
") | tee ../../gpurun_out/int8_ceiling.txt
cp ../../profiles/int8_ceiling_${ROUND}.json ../../gpurun_out/
rm -f build/hamming_tensor.o
make -j8 > /dev/null
