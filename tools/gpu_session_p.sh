#!/bin/bash
# ncu capture: sigmoid_map_kernel (K3a) and candidates_kernel (K1) inside the bench step
mkdir -p gpurun_out
NCU="ncu --clock-control none"
timeout 900 $NCU --set full --import-source on -k regex:sigmoid_map_kernel -s 3 -c 1 -o gpurun_out/r02_smap python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/p_ncu_smap.log 2>&1; echo "smap rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:candidates_kernel -s 3 -c 1 -o gpurun_out/r02_k1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/p_ncu_k1.log 2>&1; echo "k1 rc=$?"
