#!/bin/bash
mkdir -p gpurun_out
python tools/host_sweep.py > gpurun_out/host_sweep.json 2> gpurun_out/host_sweep.err; cat gpurun_out/host_sweep.err | tail -12
ISB_HOST_POOL_SLOTS=2 python tools/host_sweep.py 4 8 15 > gpurun_out/host_sweep_s2.json 2> gpurun_out/host_sweep_s2.err; echo "slots=2"; tail -3 gpurun_out/host_sweep_s2.err
ISB_HOST_POOL_SLOTS=3 python tools/host_sweep.py 4 8 15 > gpurun_out/host_sweep_s3.json 2> gpurun_out/host_sweep_s3.err; echo "slots=3"; tail -3 gpurun_out/host_sweep_s3.err
python tools/bench_tc_insitu.py > gpurun_out/split_on.json 2>&1; cat gpurun_out/split_on.json
ISB_NO_SPLIT=1 python tools/bench_tc_insitu.py > gpurun_out/split_off.json 2>&1; cat gpurun_out/split_off.json
python tools/bench_maskrcnn.py > gpurun_out/maskrcnn_rows.json 2> gpurun_out/maskrcnn_rows.err; grep -A1 "roi_align" gpurun_out/maskrcnn_rows.json | grep us
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
