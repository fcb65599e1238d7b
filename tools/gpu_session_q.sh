#!/bin/bash
# TMA RoIAlign final: parity + chain bench + stage bench + population bench
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_maskrcnn_gpu.py -m gpu -q -x > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/q_tests.log
timeout -s KILL 600 python bench.py --no-cpu-baseline > gpurun_out/q_bench_tma.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/q_bench_tma.json",):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    st=d["maskrcnn_variant"]["stages"]
    print(f, {k[:28]: round(v["us"],1) for k,v in st.items()}, d["maskrcnn_variant"]["total_us"])
PY
timeout -s KILL 300 python tools/bench_maskrcnn.py > gpurun_out/q_stage_tma.json 2> gpurun_out/q_stage.err; echo "stage rc=$?"; grep -A1 "roi_align" gpurun_out/q_stage_tma.json | grep us
timeout -s KILL 300 python tools/bench_roi_shapes.py > gpurun_out/q_roi_shapes_tma.json 2> gpurun_out/q_rs.err; cat gpurun_out/q_roi_shapes_tma.json | tr -d '\n '; echo
ISB_ROI_NO_TMA=1 timeout -s KILL 300 python tools/bench_roi_shapes.py > gpurun_out/q_roi_shapes_rows.json 2> gpurun_out/q_rs.err; cat gpurun_out/q_roi_shapes_rows.json | tr -d '\n '; echo
