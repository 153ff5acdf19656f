#!/bin/bash
# fusion-section forks (M2F_FUSION_FORK): model-level parity, then A/B of the step
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_dropout_parity.py tests/test_gpu_configs.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-modes --no-extra --no-context --no-rooflines"
for v in 0 1 0 1; do
  ms=$(M2F_FUSION_FORK=$v timeout 120 $B 2>/dev/null | tail -1 | python -c "import sys,json; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "M2F_FUSION_FORK=$v: $ms ms"
done
