B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-modes --no-extra --no-context --no-rooflines"
for v in 1 3 4 1 3 4; do
  ms=$(M2F_GEMM_BN192=$v timeout 120 $B 2>/dev/null | tail -1 | python -c "import sys,json; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "M2F_GEMM_BN192=$v: $ms ms"
done
