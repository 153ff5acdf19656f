# The staged LayerNorm kernels (csrc/layernorm.cu: ln_fwd_staged_kernel with M2F_LN_V2=1 -- op parity green, 0.74 of the HBM
# peak at the C5 size; ln_bwd_staged_kernel added with M2F_LN_V2=2 -- never run): parity of the ops and of the model first,
# then the HBM fractions at the C5 size (roofline.hbm_kernels[0..1]) and the step time.  usage: bash tools/gpu_ln_v2.sh [level]
set -x
LVL=${1:-2}
M2F_LN_V2=$LVL timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_ops.py -k "layernorm" -x -q 2>&1 | tail -5
M2F_LN_V2=$LVL timeout 300 python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py -x -q -m gpu 2>&1 | tail -3
for v in 0 1 $LVL; do
M2F_LN_V2=$v timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); ks=(d['roofline'].get('hbm_kernels') or [{},{}]); k=ks[0]; kb=ks[1]
        print('LN_V2=$v step ms', round(d['ms_per_step'],4), 'ln_fwd us', k.get('us_per_launch'), 'frac', k.get('frac'), 'ln_bwd us', kb.get('us_per_launch'), 'frac', kb.get('frac'), d['roofline'].get('hbm_kernels_error'))
"
done
