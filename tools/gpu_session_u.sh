#!/bin/bash
# one full ncu capture of the horizontal resize pass
mkdir -p gpurun_out
timeout 600 ncu --clock-control none --set full --import-source on -k regex:prep_horizontal -s 4 -c 1 -o gpurun_out/r02_prep_h -f python tools/bench_preprocess.py > gpurun_out/u_ncu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/u_ncu.log
