#!/bin/bash
# Diagnostics session: host delivery timeline, RoIAlign A/B, tensor-core benches, then tests.
mkdir -p gpurun_out
python tools/host_trace.py > gpurun_out/host_trace.txt 2>&1; tail -75 gpurun_out/host_trace.txt | head -72
python tools/host_trace.py 6 > gpurun_out/host_trace_t6.txt 2>&1; grep -E "call|isb host\] B" gpurun_out/host_trace_t6.txt | tail -3
python tools/bench_maskrcnn.py > gpurun_out/maskrcnn_rows.json 2> gpurun_out/maskrcnn_rows.err; head -30 gpurun_out/maskrcnn_rows.json
ISB_ROI_STAGED=1 python tools/bench_maskrcnn.py > gpurun_out/maskrcnn_staged.json 2> gpurun_out/maskrcnn_staged.err; head -24 gpurun_out/maskrcnn_staged.json
timeout 300 python tools/bench_mask_logits.py > gpurun_out/mask_logits.json 2> gpurun_out/mask_logits.err; cat gpurun_out/mask_logits.json; tail -3 gpurun_out/mask_logits.err
timeout 300 python tools/bench_tc_insitu.py > gpurun_out/tc_insitu_ffma.json 2>&1; cat gpurun_out/tc_insitu_ffma.json
ISB_TC_SMAP=1 timeout 300 python tools/bench_tc_insitu.py > gpurun_out/tc_insitu_tc.json 2>&1; cat gpurun_out/tc_insitu_tc.json
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
