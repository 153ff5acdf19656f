#!/bin/bash
# round 2: same-box 1-GPU and 2-GPU step times (trimmed bench lines) + the fused-optimizer tests
mkdir -p gpurun_out
python -m pytest tests/test_gpu_model.py tests/test_gpu_exchange.py -m gpu -q -p no:cacheprovider -k "adam or update_behind or train_loop" 2>&1 | tail -3
python bench.py --no-modes --no-context --no-cpu-baseline --no-rooflines --no-extra --no-e2e --steps 30 > gpurun_out/r2_q1.json 2> gpurun_out/r2_q1.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_q1.json').read().strip().splitlines()[-1]); print('N=1 ms', d['ms_per_step'])"
for extra in "" "--no-overlap-adam"; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 --steps 30 --warmup 5 --no-modes --no-context --no-cpu-baseline --no-rooflines --no-extra --no-e2e $extra > gpurun_out/r2_q2.json 2> gpurun_out/r2_q2.err
python -c "
import json
for l in open('gpurun_out/r2_q2.json'):
    if l.startswith('{'):
        d=json.loads(l); print('N=2 $extra ms', d['ms_per_step'], 'verified', d['config'].get('exchange_verified'))"
done
