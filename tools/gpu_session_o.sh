#!/bin/bash
# K2 sort-free symmetric walk: parity, then the stage timings
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_yolact_gpu.py -m gpu -q -x > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/o_tests.log
timeout -s KILL 300 python bench.py --no-cpu-baseline --no-variants > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/o_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['stage_ms'], d['roofline']['path_frac'], d['e2e']['value'])
PY
