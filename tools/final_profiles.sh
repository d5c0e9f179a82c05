#!/bin/bash
# Refreshes everything under profiles/ that describes the CURRENT kernels, in one GPU-box call:
#   bash tools/final_profiles.sh            (writes into gpurun_out/final/)
# then, back on the CPU box:  python tools/ncu_summary.py gpurun_out/final/hamming.ncu-rep profiles/...
# Every step runs under its own timeout; a number printed under ncu is never a bench value.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/final
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $OUT/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1
timeout 200 python bench.py > $OUT/bench_c3.json 2> $OUT/bench_c3.err
timeout 200 python bench.py --impl reference > $OUT/bench_ref.json 2> $OUT/bench_ref.err
timeout 120 python bench.py --mode crosscheck --no-cpu-baseline > $OUT/bench_c3_cc.json 2>/dev/null
timeout 200 python bench.py --workload c5 --no-cpu-baseline --steps 3 > $OUT/bench_c5.json 2>/dev/null
timeout 100 python bench.py --workload c1 --no-cpu-baseline > $OUT/bench_c1.json 2>/dev/null
# launch list of the default command (shares of the step), then one full capture of the top kernel
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_launches.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:hamming_tcgen05 -c 1 -f -o $OUT/hamming \
  python bench.py --steps 1 --warmup 0 --no-cpu-baseline > $OUT/ncu_full.log 2>&1
ls -la $OUT
