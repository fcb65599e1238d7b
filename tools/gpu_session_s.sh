#!/bin/bash
# end-of-round evidence on the final code: full GPU suite, bench line + reference arm, stage timings
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q -x > gpurun_out/s_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/s_tests.log
timeout -s KILL 600 python bench.py > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err; echo "bench rc=$?"
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s_bench_reference.json 2> gpurun_out/s_bench_ref.err; echo "reference rc=$?"
timeout -s KILL 300 python tools/bench_maskrcnn.py > gpurun_out/s_stage.json 2> gpurun_out/s_stage.err; echo "stage rc=$?"
timeout -s KILL 300 python tools/bench_verseg.py > gpurun_out/s_verseg.json 2> gpurun_out/s_verseg.err; echo "verseg rc=$?"
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
