#!/bin/bash
# Run under `gpurun --gpus N` with N=... in the environment: the default bench line (all sub-records) at N GPUs + the stage-by-stage e2e diagnosis
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-8}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 \
  bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/multi_n${N}_default.json 2> gpurun_out/multi_n${N}_default.err
echo "default exit $?" > gpurun_out/multi_n${N}_status.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29562 \
  tools/e2e_diag.py > gpurun_out/multi_n${N}_e2e_diag.txt 2> gpurun_out/multi_n${N}_e2e_diag.err
echo "diag exit $?" >> gpurun_out/multi_n${N}_status.txt
