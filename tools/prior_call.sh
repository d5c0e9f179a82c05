#!/bin/bash
# kNN-2 with predicted thresholds: parity tests, then A/B on one box (off / margins 1..4), C3 and C5
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_hamming_gpu.py -q -m gpu -x > gpurun_out/p_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/p_pytest.log
SSP_B200_KNN_PRIOR=0 C5=1 OUT=../../gpurun_out/p_sweep_off.txt timeout 600 bash tools/tc_sweep.sh "" > /dev/null 2>&1
C5=1 OUT=../../gpurun_out/p_sweep_on.txt timeout 1500 bash tools/tc_sweep.sh "" "-DTC_PRIOR_MARGIN=1" "-DTC_PRIOR_MARGIN=3" "-DTC_PRIOR_MARGIN=4" > /dev/null 2>&1
cat gpurun_out/p_sweep_off.txt gpurun_out/p_sweep_on.txt
