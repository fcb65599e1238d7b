#!/bin/bash
# K3a shuffle / division trims + deferred peer stores: parity (1 GPU), 2-GPU bench with the fused exchange checked against NCCL
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_yolact_gpu.py -m gpu -q -x > gpurun_out/r_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r_tests.log
timeout -s KILL 300 python bench.py --no-cpu-baseline --no-variants > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['stage_ms'], d['roofline']['path_frac'], d['e2e']['value'], d.get('from_logits_variant',{}).get('candidates_ms'))
PY
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r_n2.json 2> gpurun_out/r_n2.err; echo "n2 rc=$?"; tail -2 gpurun_out/r_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r_n2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['stage_ms'], d['config']['count_exchange'])
PY
