#!/bin/bash
# round 2, late: the 2-GPU tests and the same-box 1- / 2-GPU step times after the tile / attention / LayerNorm changes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -m gpu -q -p no:cacheprovider 2>&1 | tail -4 | tee gpurun_out/r2_pytest_gpu_2gpu.log
python bench.py --no-modes --no-context --no-cpu-baseline --no-rooflines --no-extra --no-e2e --steps 30 > gpurun_out/r2_q1.json 2> gpurun_out/r2_q1.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_q1.json').read().strip().splitlines()[-1]); print('N=1 ms', d['ms_per_step'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 --steps 30 --warmup 5 --no-modes --no-context --no-cpu-baseline --no-rooflines --no-extra > gpurun_out/r2_bench_2gpu.json 2> gpurun_out/r2_q2.err
python -c "
import json
for l in open('gpurun_out/r2_bench_2gpu.json'):
    if l.startswith('{'):
        d=json.loads(l); print('N=2 ms', d['ms_per_step'], 'value', d['value'], 'verified', d['config'].get('exchange_verified'), 'e2e', d.get('e2e',{}).get('value'))"
