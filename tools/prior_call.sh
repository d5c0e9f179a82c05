#!/bin/bash
# kNN-2 with predicted thresholds: parity tests, then A/B/A/B on one box (off / on), C3 and C5; margins 1 and 3
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_hamming_gpu.py -q -m gpu -k "predictions or c5_sized or tie_stress or ragged or golden" > gpurun_out/p_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/p_pytest.log
line() {  # label, workload
  python bench.py --steps 3 --warmup 2 --workload $2 --no-cpu-baseline --no-sub 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$1 $2 kernel_ms', round(d['roofline']['kernel_ms_per_launch'],2), 'ms_per_step', round(d['ms_per_step'],2), 'sm_mhz', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'matches', d['matches_per_step'], 'redone', d.get('redone_rows_per_step'))"
}
: > gpurun_out/p_ab.txt
for rep in 1 2; do
  for wl in c3 c5; do
    SSP_B200_KNN_PRIOR=0 line "prior=off" $wl | tee -a gpurun_out/p_ab.txt
    SSP_B200_KNN_PRIOR=1 line "prior=on(margin2)" $wl | tee -a gpurun_out/p_ab.txt
  done
done
cd opencv-starry-sky-panorama-stitcher_b200/csrc
for m in 4 6 8; do
  rm -f build/hamming_tensor.o; make EXTRA="-DTC_PRIOR_MARGIN=$m" -j8 > /dev/null 2>&1
  (cd ../.. && for wl in c3 c5; do line "prior=on(margin$m)" $wl | tee -a gpurun_out/p_ab.txt; done)
done
rm -f build/hamming_tensor.o; make -j8 > /dev/null 2>&1
