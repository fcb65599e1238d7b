#!/bin/bash
# final K1L: full GPU suite, then one full ncu capture of the kernel
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q -x > gpurun_out/w_tests.log 2>&1; echo "gpu tests rc=$?"; tail -2 gpurun_out/w_tests.log
timeout 900 ncu --clock-control none --set full --import-source on -k regex:candidates_logits -s 3 -c 1 -o gpurun_out/r02_k1l_final -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/w_ncu.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/w_ncu.log
