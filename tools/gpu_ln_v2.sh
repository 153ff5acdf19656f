# The staged LayerNorm forward (csrc/layernorm.cu: ln_fwd_staged_kernel, M2F_LN_V2=1; written without a GPU at hand):
# parity of the op and of the model first, then its HBM fraction at the C5 size (roofline.hbm_kernels[0]) and the step time.
set -x
M2F_LN_V2=1 timeout 300 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_ops.py -k "layernorm" -x -q 2>&1 | tail -5
M2F_LN_V2=1 timeout 300 python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py -x -q -m gpu 2>&1 | tail -3
for v in 0 1; do
M2F_LN_V2=$v timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); k=(d['roofline'].get('hbm_kernels') or [{}])[0]
        print('LN_V2=$v step ms', round(d['ms_per_step'],4), 'ln_fwd us', k.get('us_per_launch'), 'frac', k.get('frac'), d['roofline'].get('hbm_kernels_error'))
"
done
