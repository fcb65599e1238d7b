#!/bin/bash
# no-alloc entry point alone in a fresh process, then the whole GPU suite (it runs last there), then the bench lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/j_zz_alone.log 2>&1; echo "zz alone rc=$?"
tail -3 gpurun_out/j_zz_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/j_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -5 gpurun_out/j_gpu_tests.log
timeout 600 python bench.py > gpurun_out/bench_r02_j.json 2> gpurun_out/bench_r02_j.err; echo "bench rc=$?"
cat gpurun_out/bench_r02_j.json | head -c 3000
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r02_j_reference.json 2>> gpurun_out/bench_r02_j.err; echo "ref rc=$?"
cat gpurun_out/bench_r02_j_reference.json | head -c 800
