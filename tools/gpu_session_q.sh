#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --clock-control none --set full --import-source on -k regex:rpn_filter_fused -s 2 -c 1 -o gpurun_out/r02_rpn_fused python tools/prof_rpn.py > gpurun_out/q_ncu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/q_ncu.log
