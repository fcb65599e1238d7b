#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_maskrcnn_gpu.py -m gpu -q -x -k "roi_align and not rows_kernel" 2>&1 | tail -2
timeout -s KILL 300 python tools/bench_roi_shapes.py 2> gpurun_out/q_rs.err | tr -d '\n '; echo
