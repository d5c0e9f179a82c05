#!/bin/bash
# round 2, GPU call 5: FMA-bound filter of the polygon compare (tests + A/B), sparse reverse launch counters
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_polygon_gpu.py -q -m gpu > gpurun_out/c5_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c5_pytest.log
C4=1 timeout 900 bash tools/poly_sweep.sh "" "-DPQ_FMA_BOUND=0" "-DPQ_UNROLL=8" > gpurun_out/c5_poly_sweep.log 2>&1
cp gpurun_out/poly_sweep.txt gpurun_out/c5_poly_sweep.txt
# counters of the REVERSE launch of the two-pass cross-check (the --set full capture came back incomplete twice):
# a short metric list needs few replay passes
timeout 400 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg \
  --clock-control none -k regex:hamming_tcgen05 -s 1 -c 1 --csv --log-file gpurun_out/c5_reverse_launch.csv \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-sub --mode crosscheck > gpurun_out/c5_ncu_rev.log 2>&1
