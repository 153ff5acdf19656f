#!/bin/bash
# 2 GPUs: DP tests with the update behind the exchange (default), then A/B on the bench step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -m gpu -q --timeout 600 -p no:cacheprovider -rA 2>&1 | grep -v "^$" | tail -30 > gpurun_out/r2_tail2_pytest.log
tail -12 gpurun_out/r2_tail2_pytest.log
run() { name=$1; shift; extra=""; if [ "$1" = "--" ]; then shift; extra="$1"; shift; fi
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 30 --warmup 5 --no-e2e --no-modes --no-extra --no-context --no-rooflines --no-cpu-baseline $extra > gpurun_out/dp2_$name.json 2> gpurun_out/dp2_$name.err
  python - gpurun_out/dp2_$name.json $name <<'PY'
import json, sys
ok = False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); c = d['config']; print('DP2', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value']), 'verified', c.get('exchange_verified'), 'sync', c.get('replicas_in_sync_after_run'), c.get('optimizer_overlap', '')[:40]); ok = True
if not ok: print('DP2', sys.argv[2], 'FAILED'); print(open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
}
run tail M2F_NOP=1
run notail -- "--no-overlap-adam" M2F_NOP=1
run tail_b32 -- "--bucket-mb 32" M2F_NOP=1
run tail_b8 -- "--bucket-mb 8" M2F_NOP=1
run tail_again M2F_NOP=1
python bench.py --steps 30 --warmup 5 --no-e2e --no-modes --no-extra --no-context --no-rooflines --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N1 ms', d['ms_per_step'])"
