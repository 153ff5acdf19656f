#!/bin/bash
# round 2, 2 GPUs: what the data-parallel tax (4.21 vs 3.74 ms) responds to -- exchange CTA shape / count, bucket size, NCCL
mkdir -p gpurun_out
run() { name=$1; shift; extra=""; if [ "$1" = "--" ]; then shift; extra="$1"; shift; fi
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 30 --warmup 5 --no-e2e --no-modes --no-extra --no-context --no-rooflines --no-cpu-baseline $extra > gpurun_out/dp2_$name.json 2> gpurun_out/dp2_$name.err
  python - gpurun_out/dp2_$name.json $name <<'PY'
import json, sys
ok = False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); print('DP2', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value']), 'verified', d['config'].get('exchange_verified'), 'calls', d['config'].get('exchange_calls_per_step')); ok = True
if not ok: print('DP2', sys.argv[2], 'FAILED')
PY
}
run default M2F_NOP=1
run shape128x16 M2F_EXCHANGE_SHAPE=128,16
run shape128x16_c32 M2F_EXCHANGE_SHAPE=128,16 M2F_EXCHANGE_CTAS=32
run shape256x16 M2F_EXCHANGE_SHAPE=256,16
run ctas8 M2F_EXCHANGE_CTAS=8
run bucket32 -- "--bucket-mb 32" M2F_NOP=1
run bucket64 -- "--bucket-mb 64" M2F_NOP=1
run bucket8 -- "--bucket-mb 8" M2F_NOP=1
run nccl -- "--exchange nccl" M2F_NOP=1
run default_again M2F_NOP=1
python bench.py --steps 30 --warmup 5 --no-e2e --no-modes --no-extra --no-context --no-rooflines --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N1 ms', d['ms_per_step'])"
