#!/bin/bash
# final state with the PDL chain: full GPU suite on GPU 0, bench N=1 (full line) and N=2
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r_tests.log
timeout -s KILL 600 python bench.py > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['path_frac'], d['e2e']['value'], d['gpu_launches'])
PY
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r_n2.json 2> gpurun_out/r_n2.err; echo "n2 rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r_n2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['config']['count_exchange'][:40])
PY
timeout 900 ncu --clock-control none --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-variants > gpurun_out/r_ncu_bench.log 2>&1; echo "launch list rc=$?"
