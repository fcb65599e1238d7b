#!/bin/bash
# tests (no-alloc file alone first), then the round-2 ncu evidence: launch lists + full captures of the changed kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/k_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -2 gpurun_out/k_zz_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/k_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/k_gpu_tests.log
NCU="ncu --clock-control none"
# launch lists (per-launch durations; cold-cache, serialised)
timeout 900 $NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/k_ncu_bench.log 2>&1; echo "launch list bench rc=$?"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 1500 --csv --log-file gpurun_out/r02_maskrcnn_launches.csv python tools/bench_maskrcnn.py > gpurun_out/k_ncu_mr.log 2>&1; echo "launch list maskrcnn rc=$?"
# full captures
timeout 900 $NCU --set full --import-source on -k regex:ms_roi_align -s 2 -c 2 -o gpurun_out/r02_roi_rows python tools/prof_roi.py > gpurun_out/k_ncu_roi.log 2>&1; echo "roi rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:mask_logits_batch_tc -s 1 -c 1 -o gpurun_out/r02_mask_tc python tools/bench_mask_logits.py > gpurun_out/k_ncu_tc.log 2>&1; echo "tc rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:mask_compose_kernel -s 3 -c 1 -o gpurun_out/r02_compose python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/k_ncu_compose.log 2>&1; echo "compose rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:fast_nms_kernel -s 3 -c 1 -o gpurun_out/r02_fast_nms python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/k_ncu_fnms.log 2>&1; echo "fnms rc=$?"
# sanitizer (closed on this pool in earlier calls: record the answer)
timeout 300 compute-sanitizer --tool memcheck python -m pytest tests -m gpu -q -k "tiny" > gpurun_out/k_sanitizer.log 2>&1; echo "sanitizer rc=$?"; tail -3 gpurun_out/k_sanitizer.log
ls -la gpurun_out | head -40
