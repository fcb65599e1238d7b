#!/bin/bash
# ncu captures: fast_nms_kernel (sort-free), candidates_logits_kernel, VerSeg compose
mkdir -p gpurun_out
NCU="ncu --clock-control none"
timeout 900 $NCU --set full --import-source on -k regex:fast_nms_kernel -s 3 -c 1 -o gpurun_out/r02_fast_nms_v4 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/p_ncu_fnms.log 2>&1; echo "fnms rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:candidates_logits_kernel -s 2 -c 1 -o gpurun_out/r02_k1l python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/p_ncu_k1l.log 2>&1; echo "k1l rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:mask_compose_kernel -s 2 -c 1 -o gpurun_out/r02_verseg_compose python tools/bench_verseg.py > gpurun_out/p_ncu_vs.log 2>&1; echo "verseg rc=$?"
