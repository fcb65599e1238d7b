#!/bin/bash
# 2-GPU session: scaling bench with the fused NVLink count exchange and, for comparison, the NCCL all-gather.
mkdir -p gpurun_out
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --no-variants > gpurun_out/bench_n${N}_fused.json 2> gpurun_out/bench_n${N}_fused.err; echo "fused rc=$?"
tail -3 gpurun_out/bench_n${N}_fused.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 40 --warmup 5 --no-variants --nccl-counts > gpurun_out/bench_n${N}_nccl.json 2> gpurun_out/bench_n${N}_nccl.err; echo "nccl rc=$?"
python bench.py --gpus 1 --steps 40 --warmup 5 --no-variants --no-cpu-baseline > gpurun_out/bench_n1_ref.json 2> gpurun_out/bench_n1_ref.err
python - <<PY
import json
for f in ['bench_n${N}_fused','bench_n${N}_nccl','bench_n1_ref']:
    try:
        d=json.loads(open('gpurun_out/'+f+'.json').read().strip().splitlines()[-1])
        print(f, 'value', round(d['value']), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), d['config'].get('count_exchange','')[:40])
    except Exception as e:
        print(f, 'ERR', e)
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/pcie_bw.py > gpurun_out/pcie_bw_n${N}.json 2> gpurun_out/pcie_bw_n${N}.err; cat gpurun_out/pcie_bw_n${N}.json
python tools/pcie_bw.py > gpurun_out/pcie_bw_n1.json 2>/dev/null; cat gpurun_out/pcie_bw_n1.json
nproc
