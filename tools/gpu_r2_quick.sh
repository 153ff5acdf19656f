#!/bin/bash
# quick check after a kernel change: attention / model parity tests, then the step time twice
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -k "attention or layernorm" -p no:cacheprovider 2>&1 | tail -4
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_dropout_parity.py -m gpu -q -p no:cacheprovider 2>&1 | tail -3
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-modes --no-extra --no-context --no-rooflines"
for v in 1 2; do
  ms=$(timeout 120 $B 2>/dev/null | tail -1 | python -c "import sys,json; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "step: $ms ms"
done
