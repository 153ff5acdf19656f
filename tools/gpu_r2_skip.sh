#!/bin/bash
# Upper bounds: the train step (C2, bf16x2f) with whole kernel classes left out of the plan (M2F_DEBUG_SKIP, results garbage, timing valid)
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-modes --no-extra --no-context --no-rooflines"
for skip in "" m2f_ln_fwd m2f_ln_bwd m2f_ln_fwd,m2f_ln_bwd m2f_attn_fwd m2f_attn_bwd m2f_attn_fwd,m2f_attn_bwd m2f_ln,m2f_attn,m2f_split,m2f_relu; do
  ms=$(M2F_DEBUG_SKIP="$skip" timeout 120 $B 2>/dev/null | tail -1 | python -c "import sys,json; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "skip [$skip]: $ms ms"
done
