#!/bin/bash
# more upper bounds (M2F_DEBUG_SKIP): weight gradients (off the critical path), the forward / backward GEMMs by call text
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-modes --no-extra --no-context --no-rooflines"
for skip in "" wgrad "c.gemm(d" "c.gemm(cur_pl,c.gemm(o,c.gemm(x1p,c.gemm(h" "c.gemm(t_cur,c.gemm(xo,c.gemm(u,c.gemm(a" "c.gemm"; do
  ms=$(M2F_DEBUG_SKIP="$skip" timeout 120 $B 2>/dev/null | tail -1 | python -c "import sys,json; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "skip [$skip]: $ms ms"
done
