#!/bin/bash
# 4-GPU bench line exactly as the driver launches it (final code of the round)
mkdir -p gpurun_out
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 4 --steps 50 --warmup 5 > gpurun_out/n4b_bench.json 2> gpurun_out/n4b_bench.err; echo "n4 rc=$?"; head -c 300 gpurun_out/n4b_bench.json; tail -2 gpurun_out/n4b_bench.err
