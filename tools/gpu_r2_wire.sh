#!/bin/bash
# round 2, 2 GPUs: bf16 wire format of the gradient exchange -- tests, then fp32 vs bf16 wire on the bench step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -m gpu -q --timeout 600 -p no:cacheprovider -rA 2>&1 | grep -v "^$" | tail -30 > gpurun_out/r2_wire_pytest.log
tail -14 gpurun_out/r2_wire_pytest.log
run() { name=$1; shift; extra=""; if [ "$1" = "--" ]; then shift; extra="$1"; shift; fi
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 30 --warmup 5 --no-e2e --no-modes --no-extra --no-context --no-rooflines --no-cpu-baseline $extra > gpurun_out/dp2_$name.json 2> gpurun_out/dp2_$name.err
  python - gpurun_out/dp2_$name.json $name <<'PY'
import json, sys
ok = False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); c = d['config']; print('DP2', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value']), 'wire', c.get('exchange_wire'), 'verified', c.get('exchange_verified'), c.get('exchange_check', {}) and {k: c['exchange_check'][k] for k in ('rel_err_2d_weights', 'rel_err_1d_params')}, 'calls', c.get('exchange_calls_per_step')); ok = True
if not ok: print('DP2', sys.argv[2], 'FAILED'); print(open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
}
run wire_bf16 M2F_NOP=1
run wire_fp32 -- "--wire fp32" M2F_NOP=1
run wire_bf16_b32 -- "--bucket-mb 32" M2F_NOP=1
run wire_bf16_b8 -- "--bucket-mb 8" M2F_NOP=1
run wire_bf16_again M2F_NOP=1
