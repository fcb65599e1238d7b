#!/bin/bash
mkdir -p gpurun_out
python tools/host_trace.py > gpurun_out/host_trace.txt 2>&1; grep -E "call|isb host\] B" gpurun_out/host_trace.txt | tail -4; grep "isb host\] img" gpurun_out/host_trace.txt | tail -64 | awk 'NR%4==1' 
python tools/host_trace.py 6 > gpurun_out/host_trace_t6.txt 2>&1; grep -E "call" gpurun_out/host_trace_t6.txt | tail -2
python tools/host_trace.py 0 > gpurun_out/host_trace_t0.txt 2>&1; grep -E "call" gpurun_out/host_trace_t0.txt | tail -2
python tools/bench_maskrcnn.py > gpurun_out/maskrcnn_rows.json 2> gpurun_out/maskrcnn_rows.err; grep -A1 "roi_align" gpurun_out/maskrcnn_rows.json | grep us
ISB_ROI_STAGED=1 python tools/bench_maskrcnn.py > gpurun_out/maskrcnn_staged.json 2> gpurun_out/maskrcnn_staged.err; grep -A1 "roi_align" gpurun_out/maskrcnn_staged.json | grep us
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -6 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02_b.json 2> gpurun_out/bench_r02_b.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r02_b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02_b.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',json.dumps(d['e2e'])[:700])
print('sparse',json.dumps(d['sparse_variant'])[:400])
mv=d['maskrcnn_variant']; print({k[:24]:round(v['us'],1) for k,v in mv['stages'].items()}, mv['total_us'])
PY
