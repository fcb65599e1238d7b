#!/bin/bash
# One GPU-box session: full GPU test suite, bench, sanitizer passes over the small tests, ncu captures of the new kernels.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02_a.json 2> gpurun_out/bench_r02_a.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_r02_a.json; tail -5 gpurun_out/bench_r02_a.err
SEL='tiny or stage or paste_odd or paste_chunking or roi_align_edge or two_images or count_exchange or nan_inf or assemble_masks_vs or detect_host_matches'
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 --log-file gpurun_out/sanitizer_memcheck.txt \
  python -m pytest tests -m gpu -q -x -k "$SEL" > gpurun_out/sanitizer_memcheck_pytest.log 2>&1; echo "memcheck rc=$?" | tee -a gpurun_out/sanitizer_memcheck_pytest.log
tail -3 gpurun_out/sanitizer_memcheck.txt
SEL2='tiny_end_to_end or assemble_masks_vs or paste_odd or two_images or count_exchange or test_roi_align_edge'
timeout 1200 compute-sanitizer --tool racecheck --error-exitcode 9 --log-file gpurun_out/sanitizer_racecheck.txt \
  python -m pytest tests -m gpu -q -x -k "$SEL2" > gpurun_out/sanitizer_racecheck_pytest.log 2>&1; echo "racecheck rc=$?" | tee -a gpurun_out/sanitizer_racecheck_pytest.log
tail -3 gpurun_out/sanitizer_racecheck.txt
timeout 600 compute-sanitizer --tool synccheck --error-exitcode 9 --log-file gpurun_out/sanitizer_synccheck.txt \
  python -m pytest tests -m gpu -q -x -k "$SEL2" > gpurun_out/sanitizer_synccheck_pytest.log 2>&1; echo "synccheck rc=$?" | tee -a gpurun_out/sanitizer_synccheck_pytest.log
tail -3 gpurun_out/sanitizer_synccheck.txt
