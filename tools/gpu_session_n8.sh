#!/bin/bash
# N-GPU check of the fused NVLink count exchange: bench at N = $1
N=${1:-8}
mkdir -p gpurun_out
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --no-cpu-baseline > gpurun_out/n${N}_bench_fused.json 2> gpurun_out/n${N}_bench_fused.err; echo "fused rc=$?"; head -c 1300 gpurun_out/n${N}_bench_fused.json; tail -3 gpurun_out/n${N}_bench_fused.err
