#!/bin/bash
mkdir -p gpurun_out
python tools/host_trace.py 8 > gpurun_out/host_trace_t8.txt 2>&1; grep -E "call" gpurun_out/host_trace_t8.txt | tail -2; grep "chunk" gpurun_out/host_trace_t8.txt | tail -16
python tools/bench_tc_insitu.py > gpurun_out/split_on.json 2>&1; cat gpurun_out/split_on.json
ISB_NO_SPLIT=1 python tools/bench_tc_insitu.py > gpurun_out/split_off.json 2>&1; cat gpurun_out/split_off.json
