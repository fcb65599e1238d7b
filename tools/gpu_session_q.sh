#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python bench.py --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/q_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/q_bench.json").read().strip().splitlines()[-1])
mv=d["maskrcnn_variant"]
print({k[:28]: round(v["us"],1) for k,v in mv["stages"].items()}, mv["total_us"], mv["sync_free_pipeline"])
PY
