#!/bin/bash
# 2-GPU bench line exactly as the driver launches it (final code of the round)
mkdir -p gpurun_out
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/n2b_bench.json 2> gpurun_out/n2b_bench.err; echo "n2 rc=$?"; head -c 400 gpurun_out/n2b_bench.json; tail -2 gpurun_out/n2b_bench.err
