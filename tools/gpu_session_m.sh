#!/bin/bash
# re-entry check: full GPU suite, default bench line, reference arm
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q -x > gpurun_out/m_tests.log 2>&1; echo "gpu tests rc=$?"; tail -8 gpurun_out/m_tests.log
timeout -s KILL 600 python bench.py > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; echo "bench rc=$?"; head -c 1500 gpurun_out/m_bench.json
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/m_smoke.log
