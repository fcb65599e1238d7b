#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_maskrcnn_gpu.py -m gpu -q -x 2>&1 | tail -3
timeout -s KILL 300 python tools/bench_maskrcnn.py > gpurun_out/q_stage.json 2> gpurun_out/q_stage.err; python -c "
import json; d=json.load(open('gpurun_out/q_stage.json'))
print({k:round(v['us'],1) for k,v in d.items() if isinstance(v,dict) and 'us' in v})"
