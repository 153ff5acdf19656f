# 1-GPU A/B of environment switches on the bench step (device-resident inputs only)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model.py -x -q -m gpu 2>&1 | tail -2
ab() { name=$1; shift; env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err; python - gpurun_out/ab_$name.json $name <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); print('AB', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value']))
PY
}
ab default M2F_NOP=1
ab prio0 M2F_STREAM_PRIO=0
ab default_again M2F_NOP=1
ab cta2_0 M2F_GEMM_CTA2=0
ab cta2_1 M2F_GEMM_CTA2=1
ab pdl0 M2F_PDL=0
