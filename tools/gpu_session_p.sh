#!/bin/bash
# ncu capture of fast_nms_kernel after the counting-sort change
mkdir -p gpurun_out
NCU="ncu --clock-control none"
timeout 900 $NCU --set full --import-source on -k regex:fast_nms_kernel -s 3 -c 1 -o gpurun_out/r02_fast_nms_v3 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/p_ncu_fnms.log 2>&1; echo "fnms rc=$?"
