#!/bin/bash
# round 2, GPU call 7: full validation of the tree -- whole GPU suite, smoke, default bench line, reference arm, stage bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -q -m gpu ) > gpurun_out/c7_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c7_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c7_smoke.log 2>&1
( time timeout 900 python bench.py --steps 20 --warmup 3 ) > gpurun_out/c7_bench_default.json 2> gpurun_out/c7_bench_default.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/c7_bench_reference.json 2> gpurun_out/c7_bench_reference.err
timeout 300 python bench.py --workload c1real --steps 3 --warmup 1 --ransac device > gpurun_out/c7_bench_c1real_device.json 2> gpurun_out/c7_bench_c1real_device.err
timeout 300 python bench.py --workload c1 --steps 20 --warmup 3 --no-cpu-baseline --mode crosscheck > gpurun_out/c7_bench_c1_cross.json 2> gpurun_out/c7_bench_c1_cross.err
