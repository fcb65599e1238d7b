#!/bin/bash
# compose kBig variant: parity, headline bench (must not regress), VerSeg timings + ncu
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_yolact_gpu.py -m gpu -q -x > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/o_tests.log
timeout -s KILL 300 python bench.py --no-cpu-baseline --no-variants > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/o_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['stage_ms'], d['roofline']['path_frac'], d['e2e']['value'])
PY
for i in 1 2; do timeout -s KILL 300 python tools/bench_verseg.py > gpurun_out/o_verseg$i.json 2> gpurun_out/o_verseg.err; echo "verseg rc=$?"; grep -E "assemble_masks|device_only" gpurun_out/o_verseg$i.json; done
timeout 900 ncu --clock-control none --set full --import-source on -k regex:mask_compose_kernel -s 2 -c 1 -o gpurun_out/r02_verseg_compose_v2 python tools/bench_verseg.py > gpurun_out/p_ncu_vs.log 2>&1; echo "ncu verseg rc=$?"
