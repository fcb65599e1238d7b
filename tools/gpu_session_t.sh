#!/bin/bash
# horizontal resize pass: parity (bit-equal to Pillow) + timing, 192 vs 128 columns per block, per-kernel times
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_preprocess_gpu.py -m gpu -q -x > gpurun_out/t_tests.log 2>&1; echo "prep tests rc=$?"; tail -1 gpurun_out/t_tests.log
timeout -s KILL 300 python tools/bench_preprocess.py > gpurun_out/t_prep.json 2> gpurun_out/t_prep.err; echo "prep rc=$?"; grep -A1 550x550 gpurun_out/t_prep.json
ISB_PREP_T128=1 timeout -s KILL 300 python tools/bench_preprocess.py > gpurun_out/t_prep128.json 2> gpurun_out/t_prep128.err; echo "prep128 rc=$?"; grep -A1 550x550 gpurun_out/t_prep128.json
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/t_launches.csv python tools/bench_preprocess.py > /dev/null 2>&1; echo "ncu rc=$?"
grep -c prep_ gpurun_out/t_launches.csv
