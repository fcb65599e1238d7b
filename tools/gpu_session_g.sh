#!/bin/bash
mkdir -p gpurun_out
python tools/host_sweep.py 6 8 10 15 > gpurun_out/host_sweep_2up.json 2> gpurun_out/host_sweep_2up.err; echo two_up; tail -4 gpurun_out/host_sweep_2up.err
ISB_HOST_ONE_UP=1 python tools/host_sweep.py 8 > gpurun_out/host_sweep_1up.json 2> gpurun_out/host_sweep_1up.err; echo one_up; tail -1 gpurun_out/host_sweep_1up.err
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02_c.json 2> gpurun_out/bench_r02_c.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r02_c.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02_c.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',json.dumps(d['e2e'])[:300])
print('roofline', json.dumps(d['roofline'])[:500])
mv=d['maskrcnn_variant']; print({k[:24]:round(v['us'],1) for k,v in mv['stages'].items()}, mv['total_us'])
PY
