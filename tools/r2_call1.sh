#!/bin/bash
# round 2, GPU call 1: parity of the fused cross-check / compacted polygon compare + first timings
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/c1_gpu.txt 2>&1
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/c1_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c1_pytest.log
for form in fused sparse; do
  for wl in c3 c5; do
    SSP_B200_CROSSCHECK=$form timeout 600 python bench.py --workload $wl --mode crosscheck --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/c1_bench_${wl}_cross_${form}.json 2> gpurun_out/c1_bench_${wl}_cross_${form}.err
  done
done
timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c1_bench_c3_knn.json 2> gpurun_out/c1_bench_c3_knn.err
for k in queue legacy; do
  SSP_B200_POLY_COMPARE=$k timeout 600 python bench.py --workload c4cap --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/c1_bench_c4cap_${k}.json 2> gpurun_out/c1_bench_c4cap_${k}.err
  SSP_B200_POLY_COMPARE=$k timeout 600 python bench.py --workload c4 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/c1_bench_c4_${k}.json 2> gpurun_out/c1_bench_c4_${k}.err
done
# the new bench (default invocation with sub-records), if present
if [ -f bench_new.py ]; then
  timeout 900 python bench_new.py --steps 5 --warmup 3 > gpurun_out/c1_benchnew_default.json 2> gpurun_out/c1_benchnew_default.err
fi
