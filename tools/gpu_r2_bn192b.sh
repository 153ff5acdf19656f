#!/bin/bash
# 192-wide pair tiles in the backward (MN-major B through the 64-byte swizzle): op tests, model parity, A/B (bit 1 dgrad, bit 2 wgrad)
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -k "gemm" -p no:cacheprovider 2>&1 | tail -5
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_dropout_parity.py -m gpu -q -k "meld_base" -p no:cacheprovider 2>&1 | tail -3
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-modes --no-extra --no-context --no-rooflines"
for v in 0 1 2 3 0 1 2 3; do
  ms=$(M2F_GEMM_BN192_BWD=$v timeout 120 $B 2>/dev/null | tail -1 | python -c "import sys,json; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "M2F_GEMM_BN192_BWD=$v: $ms ms"
done
