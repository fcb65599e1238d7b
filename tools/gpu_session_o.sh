#!/bin/bash
# PDL between the sigmoid-map and the compose kernel: parity, then A/B of the headline step
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_yolact_gpu.py -m gpu -q -x > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/o_tests.log
for i in 1 2; do
timeout -s KILL 300 python bench.py --no-cpu-baseline --no-variants > gpurun_out/o_bench_pdl.json 2> gpurun_out/o_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/o_bench_pdl.json').read().strip().splitlines()[-1]); print('pdl   ', d['value'], d['ms_per_step'])"
ISB_NO_PDL=1 timeout -s KILL 300 python bench.py --no-cpu-baseline --no-variants > gpurun_out/o_bench_nopdl.json 2> gpurun_out/o_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/o_bench_nopdl.json').read().strip().splitlines()[-1]); print('no pdl', d['value'], d['ms_per_step'])"
done
