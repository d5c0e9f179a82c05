#!/bin/bash
# Refreshes everything under profiles/ that describes the CURRENT kernels, in one GPU-box call
# (writes into gpurun_out/final/; summarise back on the CPU box with tools/ncu_summary.py and
# tools/launch_share.py).  Every step runs under its own timeout; a number printed under ncu is
# never a bench value.  Multi-GPU lines: tools/multi_gpu_bench.sh under `gpurun --gpus N`.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/final
mkdir -p $OUT
( time timeout 2400 python -m pytest tests -q -m gpu ) > $OUT/pytest_gpu.log 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1
( time timeout 900 python bench.py --steps 20 --warmup 3 ) > $OUT/bench_default.json 2> $OUT/bench_default.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err
for r in opencv device; do
  timeout 600 python bench.py --workload c1real --steps 3 --warmup 1 --ransac $r > $OUT/bench_c1real_$r.json 2> $OUT/bench_c1real_$r.err
done
timeout 600 bash tools/int8_ceiling.sh > $OUT/int8_ceiling.log 2>&1
# launch lists (shares of a step)
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_c3_knn.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub > $OUT/ncu_l1.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_c3_cross.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub --mode crosscheck > $OUT/ncu_l2.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_c4.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --workload c4 > $OUT/ncu_l3.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_c1real.csv \
  python bench.py --steps 1 --warmup 1 --workload c1real --ransac device > $OUT/ncu_l4.log 2>&1
# full captures of the dominant kernels
timeout 500 ncu --set full --clock-control none --import-source on -k regex:hamming_tcgen05 -c 1 -f -o $OUT/hamming_knn \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub > $OUT/ncu_f1.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:hamming_tcgen05 -c 1 -f -o $OUT/hamming_cross_forward \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub --mode crosscheck > $OUT/ncu_f2.log 2>&1
# the reverse (sparse) launch of the cross-check: a short metric list completes where --set full came back incomplete
timeout 400 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg \
  --clock-control none -k regex:hamming_tcgen05 -s 1 -c 1 --csv --log-file $OUT/reverse_launch.csv \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub --mode crosscheck > $OUT/ncu_rev.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:poly_screen_tcgen05 -c 1 -f -o $OUT/poly_tensor_screen \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --workload c4cap > $OUT/ncu_f4.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches_c4cap.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload c4cap > $OUT/ncu_l5.log 2>&1
ls -la $OUT
