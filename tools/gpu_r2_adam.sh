#!/bin/bash
# round 2: Adam inside the backward -- correctness (bit-identical to step()), then A/B on the step; narrow-tile levels
mkdir -p gpurun_out
python -m pytest tests/test_gpu_adam_in_backward.py tests/test_gpu_model.py tests/test_gpu_configs.py -m gpu -q --timeout 900 -p no:cacheprovider 2>&1 | tail -25 > gpurun_out/r2_adam_pytest.log
tail -6 gpurun_out/r2_adam_pytest.log
ab() { name=$1; shift; extra=""; if [ "$1" = "--" ]; then shift; extra="$1"; shift; fi; env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline --no-modes --no-extra --no-context --no-rooflines $extra > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err; python - gpurun_out/ab_$name.json $name <<'PY'
import json, sys
ok = False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); print('AB', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value'])); ok = True
if not ok: print('AB', sys.argv[2], 'FAILED')
PY
}
ab inbwd M2F_NOP=1
ab step -- "--optimizer step" M2F_NOP=1
ab inbwd_narrow2 M2F_NARROW=2
ab inbwd_narrow3 M2F_NARROW=3
ab inbwd_narrow0 M2F_NARROW=0
ab inbwd_again M2F_NOP=1
tail -3 gpurun_out/ab_inbwd.err
