#!/bin/bash
# one full ncu capture of each resize pass (550 x 550 case: launches after the 544 series)
mkdir -p gpurun_out
timeout 600 ncu --clock-control none --set full --import-source on -k regex:prep_ -s 60 -c 2 -o gpurun_out/r02_prep_hv -f python tools/bench_preprocess.py > gpurun_out/u_ncu.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/u_ncu.log
