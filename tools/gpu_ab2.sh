set -x
mkdir -p gpurun_out
M2F_GEMM_CTA2=1 timeout 900 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -3
ab() { name=$1; shift; env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline $EXTRA > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err; python - gpurun_out/ab_$name.json $name <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); print('AB', sys.argv[2], 'ms', round(d['ms_per_step'], 4), 'value', round(d['value']), 'roof', round(d['roofline']['frac'], 3), round(d['roofline']['us_per_launch'], 2))
PY
}
ab cta2_1_bn128 M2F_GEMM_CTA2=1 M2F_GEMM_BN=128
ab cta2_1 M2F_GEMM_CTA2=1
EXTRA="--precision fp32"
ab fp32_default M2F_NOP=1
ab fp32_cta2_1 M2F_GEMM_CTA2=1
EXTRA="--batch 256"
ab b256_default M2F_NOP=1
ab b256_cta2_1 M2F_GEMM_CTA2=1
