#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --clock-control none --set full --import-source on -k regex:ms_roi_align -s 2 -c 2 -o gpurun_out/r02_roi_tma python tools/prof_roi.py > gpurun_out/q_ncu_roi.log 2>&1; echo "roi rc=$?"
