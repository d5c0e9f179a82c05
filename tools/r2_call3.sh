#!/bin/bash
# round 2, GPU call 3: re-run of the tests that failed in call 2, polygon default decision, ncu evidence
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_polygon_gpu.py tests/test_ransac_gpu.py tests/test_matcher_gpu.py -q -m gpu > gpurun_out/c3_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c3_pytest.log
timeout 400 bash tools/poly_sweep.sh "" "-DPQ_TILE_COLS=64" > gpurun_out/c3_poly_sweep.log 2>&1
cp gpurun_out/poly_sweep.txt gpurun_out/c3_poly_sweep.txt
timeout 300 python tools/n2_timing.py 8 > gpurun_out/c3_n2_timing.json 2> gpurun_out/c3_n2_timing.err
timeout 300 python bench.py --workload c1real --steps 3 --warmup 1 --ransac device > gpurun_out/c3_bench_c1real_device.json 2> gpurun_out/c3_bench_c1real_device.err
# launch lists (shares of a step) -- numbers printed under ncu are never bench values
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c3_launches_c3_knn.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub > gpurun_out/c3_ncu_l1.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c3_launches_c3_cross.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub --mode crosscheck > gpurun_out/c3_ncu_l2.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/c3_launches_c4.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --workload c4 > gpurun_out/c3_ncu_l3.log 2>&1
# full captures of the dominant kernels
timeout 500 ncu --set full --clock-control none --import-source on -k regex:hamming_tcgen05 -c 1 -f -o gpurun_out/c3_hamming_knn \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub > gpurun_out/c3_ncu_f1.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:hamming_tcgen05 -c 2 -f -o gpurun_out/c3_hamming_cross \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub --mode crosscheck > gpurun_out/c3_ncu_f2.log 2>&1
SSP_B200_CROSSCHECK=fused timeout 600 ncu --set full --clock-control none --import-source on -k regex:hamming_tcgen05 -c 1 -f -o gpurun_out/c3_hamming_fused \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub --mode crosscheck > gpurun_out/c3_ncu_f3.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:poly_compare -c 1 -f -o gpurun_out/c3_poly_queue \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload c4cap > gpurun_out/c3_ncu_f4.log 2>&1
SSP_B200_POLY_COMPARE=legacy timeout 500 ncu --set full --clock-control none --import-source on -k regex:poly_compare -c 1 -f -o gpurun_out/c3_poly_legacy \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload c4cap > gpurun_out/c3_ncu_f5.log 2>&1
ls -la gpurun_out | tail -30
