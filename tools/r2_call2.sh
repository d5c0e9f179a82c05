#!/bin/bash
# round 2, GPU call 2: full suite on the new rows (N2, N3, multi-store), polygon sweep, int8 ceiling, stage bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu > gpurun_out/c2_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c2_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c2_smoke.log 2>&1
C4=1 timeout 900 bash tools/poly_sweep.sh "" "-DPQ_TILE_COLS=64" "-DPQ_UNROLL=2" "-DPQ_TILE_COLS=256 -DPQ_UNROLL=4" "-DPQ_TILE_COLS=64 -DPQ_UNROLL=8" > gpurun_out/c2_poly_sweep.log 2>&1
timeout 600 bash tools/int8_ceiling.sh > gpurun_out/c2_int8.log 2>&1
for r in opencv device; do
  timeout 600 python bench.py --workload c1real --steps 2 --warmup 1 --ransac $r > gpurun_out/c2_bench_c1real_$r.json 2> gpurun_out/c2_bench_c1real_$r.err
done
timeout 600 python bench.py --workload c1real --impl reference --steps 2 --warmup 1 > gpurun_out/c2_bench_c1real_ref.json 2> gpurun_out/c2_bench_c1real_ref.err
