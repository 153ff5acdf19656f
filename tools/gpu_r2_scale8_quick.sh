#!/bin/bash
# round 2, late: same-box N = 1 and N = 8 step times (trimmed bench lines, no extra workloads)
mkdir -p gpurun_out
python bench.py --no-modes --no-context --no-cpu-baseline --no-rooflines --no-extra --no-e2e --steps 30 > gpurun_out/r2_q1.json 2> gpurun_out/r2_q1.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_q1.json').read().strip().splitlines()[-1]); print('N=1 ms', d['ms_per_step'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 8 --steps 30 --warmup 5 --no-modes --no-context --no-cpu-baseline --no-rooflines --no-extra > gpurun_out/r2_bench_8gpu.json 2> gpurun_out/r2_bench_8gpu.err
echo "rc=$?"; tail -c 300 gpurun_out/r2_bench_8gpu.err
python -c "
import json
for l in open('gpurun_out/r2_bench_8gpu.json'):
    if l.startswith('{'):
        d=json.loads(l); print('N=8 ms', d['ms_per_step'], 'value', d['value'], 'verified', d['config'].get('exchange_verified'), 'e2e', d.get('e2e',{}).get('value'))"
