#!/bin/bash
# one-launch rpn_filter / box_postprocess: parity on both paths, then stage timings fused vs chain
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_maskrcnn_gpu.py -m gpu -q -x > gpurun_out/l_mr_tests.log 2>&1; echo "maskrcnn tests rc=$?"; tail -15 gpurun_out/l_mr_tests.log
timeout -s KILL 300 python tools/bench_maskrcnn.py > gpurun_out/l_stage_fused.json 2> gpurun_out/l_stage_fused.err; echo "fused rc=$?"; cat gpurun_out/l_stage_fused.json | head -c 2500
ISB_RCNN_CHAIN=1 timeout -s KILL 300 python tools/bench_maskrcnn.py > gpurun_out/l_stage_chain.json 2> gpurun_out/l_stage_chain.err; echo "chain rc=$?"; cat gpurun_out/l_stage_chain.json | head -c 2500
