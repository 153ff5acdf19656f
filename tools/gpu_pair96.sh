# Build first with M2F_NVCC_DEFS="-DM2F_PAIR_RING_KB=96 -DM2F_PAIR_EPI_WARPS=4" (see DESIGN.md section 9), then run on one B200:
# GEMM op tests + the bench step with CTA pairs that share an SM two by two.  Rebuild without the variable afterwards.
timeout 25 python -m pytest tests/test_gpu_ops.py -k gemm -x -q 2>&1 | tail -2
timeout 30 python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('PAIR96 ms', round(d['ms_per_step'],4), 'roof_us', round(d['roofline']['us_per_launch'],2))
"
