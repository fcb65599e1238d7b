#!/bin/bash
# K1L with the max-only row pre-test: parity tests + the bench line's from_logits_variant
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_yolact_gpu.py -m gpu -q -x -k "logits" > gpurun_out/v_tests.log 2>&1; echo "logits tests rc=$?"; tail -1 gpurun_out/v_tests.log
timeout -s KILL 600 python bench.py --no-cpu-baseline > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/v_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['from_logits_variant'])
PY
