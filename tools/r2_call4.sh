#!/bin/bash
# round 2, GPU call 4: polygon tests on the 2-rows-per-thread compare kernel + batch measure, sweep, C4 lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_polygon_gpu.py -q -m gpu > gpurun_out/c4_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c4_pytest.log
C4=1 timeout 900 bash tools/poly_sweep.sh "" "-DPQ_UNROLL=4" "-DPQ_UNROLL=1" "-DPQ_ROWS=1 -DPQ_UNROLL=4" "-DPQ_TILE_COLS=128" > gpurun_out/c4_poly_sweep.log 2>&1
cp gpurun_out/poly_sweep.txt gpurun_out/c4_poly_sweep.txt
