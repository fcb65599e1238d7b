#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_yolact_gpu.py -m gpu -q -x -k "logits" > gpurun_out/v_tests.log 2>&1; echo "logits tests rc=$?"; tail -1 gpurun_out/v_tests.log
timeout -s KILL 600 python bench.py --no-cpu-baseline > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/v_bench.json').read().strip().splitlines()[-1])
print(d['from_logits_variant']['candidates_ms'])
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:candidates_logits -c 6 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | grep "gpu__time_duration" | tail -3
