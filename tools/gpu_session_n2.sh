#!/bin/bash
# 2-GPU check of the fused NVLink count exchange (and the NCCL fallback for A/B)
mkdir -p gpurun_out
(lscpu | head -30; nproc; numactl -H 2>/dev/null; cat /sys/fs/cgroup/cpu.max 2>/dev/null; free -g | head -2; nvidia-smi topo -m) > gpurun_out/n2_host.txt 2>&1
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/n2_bench_fused.json 2> gpurun_out/n2_bench_fused.err; echo "fused rc=$?"; head -c 1200 gpurun_out/n2_bench_fused.json; tail -3 gpurun_out/n2_bench_fused.err
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline --nccl-counts > gpurun_out/n2_bench_nccl.json 2> gpurun_out/n2_bench_nccl.err; echo "nccl rc=$?"; head -c 600 gpurun_out/n2_bench_nccl.json
