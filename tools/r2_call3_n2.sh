#!/bin/bash
# round 2, GPU call 3 (2 GPUs): the multi-GPU paths on real NCCL -- Hamming e2e chain (packed broadcast,
# exact-length gather) and the star-polygon path (star bank broadcast, measure split by image, gather)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-2}
run() {  # name, bench args...
  name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $N "$@" > gpurun_out/c3_n${N}_$name.json 2> gpurun_out/c3_n${N}_$name.err
  echo "$name exit $?" >> gpurun_out/c3_n${N}_status.txt
}
: > gpurun_out/c3_n${N}_status.txt
run default --steps 5 --warmup 3 --sub crosscheck,c4,c4cap
run c5 --workload c5 --steps 3 --warmup 3
# single GPU on the same box for the ratio
timeout 900 python bench.py --steps 5 --warmup 3 --sub crosscheck,c4,c4cap --no-cpu-baseline > gpurun_out/c3_n1_default.json 2> gpurun_out/c3_n1_default.err
echo "n1 exit $?" >> gpurun_out/c3_n${N}_status.txt
