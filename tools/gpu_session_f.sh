#!/bin/bash
mkdir -p gpurun_out
ISB_HOST_ISA=dry python tools/host_trace.py 8 > gpurun_out/host_trace_dry.txt 2>&1; echo dry; grep -E "call" gpurun_out/host_trace_dry.txt | tail -1; grep "chunk" gpurun_out/host_trace_dry.txt | tail -16 | awk 'NR%3==0'
ISB_HOST_ISA=cached python tools/host_sweep.py 4 8 15 > gpurun_out/host_sweep_cached.json 2> gpurun_out/host_sweep_cached.err; echo cached; tail -3 gpurun_out/host_sweep_cached.err
ISB_HOST_ISA=avx2 python tools/host_sweep.py 8 > gpurun_out/host_sweep_avx2.json 2> gpurun_out/host_sweep_avx2.err; echo avx2; tail -1 gpurun_out/host_sweep_avx2.err
