#!/bin/bash
# round 2: the bench line at N GPUs as the driver launches it (trimmed: no other precision modes, no eager context, no CPU arm)
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus $N --steps 20 --warmup 5 --no-modes --no-context --no-cpu-baseline --no-rooflines > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
echo "rc=$?"; tail -c 800 gpurun_out/r2_bench_${N}gpu.err
python - $N <<'PY'
import json, sys
for l in open(f'gpurun_out/r2_bench_{sys.argv[1]}gpu.json'):
    if l.startswith('{'):
        d = json.loads(l)
        print('N', d['n_gpus'], 'ms', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'] and d['e2e']['value'])
        print(json.dumps({k: v for k, v in d['config'].items() if 'exchange' in k or 'sync' in k}))
        print(json.dumps(d.get('extra_workloads')))
PY
