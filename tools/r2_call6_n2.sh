#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 tools/e2e_diag.py > gpurun_out/c6_e2e_diag.txt 2> gpurun_out/c6_e2e_diag.err
grep '^{' gpurun_out/c6_e2e_diag.txt
