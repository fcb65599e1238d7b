#!/bin/bash
# end-of-round evidence: full GPU suite, bench line + reference arm, ncu launch lists of the bench and of the Mask R-CNN stages,
# ncu of the roofline kernel and of the TMA RoIAlign
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q -x > gpurun_out/s_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/s_tests.log
timeout -s KILL 600 python bench.py > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err; echo "bench rc=$?"
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s_bench_reference.json 2> gpurun_out/s_bench_ref.err; echo "reference rc=$?"
timeout -s KILL 300 python tools/bench_maskrcnn.py > gpurun_out/s_stage.json 2> gpurun_out/s_stage.err; echo "stage rc=$?"
timeout -s KILL 300 python tools/bench_verseg.py > gpurun_out/s_verseg.json 2> gpurun_out/s_verseg.err; echo "verseg rc=$?"
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/s_ncu_bench.log 2>&1; echo "launch list bench rc=$?"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 1500 --csv --log-file gpurun_out/r02_maskrcnn_launches.csv python tools/bench_maskrcnn.py > gpurun_out/s_ncu_mr.log 2>&1; echo "launch list maskrcnn rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:mask_compose_kernel -s 3 -c 1 -o gpurun_out/r02_compose python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/s_ncu_compose.log 2>&1; echo "compose rc=$?"
timeout 900 $NCU --set full --import-source on -k regex:ms_roi_align -s 2 -c 2 -o gpurun_out/r02_roi_final python tools/prof_roi.py > gpurun_out/s_ncu_roi.log 2>&1; echo "roi rc=$?"
