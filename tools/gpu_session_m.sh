#!/bin/bash
# full check of the committed state: GPU suite, default bench line, reference arm, smoke
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q -x > gpurun_out/m_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/m_tests.log
timeout -s KILL 600 python bench.py > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; echo "bench rc=$?"; head -c 400 gpurun_out/m_bench.json; echo
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/m_bench_reference.json 2> gpurun_out/m_ref.err; echo "reference rc=$?"
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/m_smoke.log
