#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_ops.py -m gpu -q -k "layernorm" --timeout 600 -p no:cacheprovider 2>&1 | tail -4
python - <<'PY'
import importlib, json, sys, torch
sys.path.insert(0, '.')
import bench
pkg = importlib.import_module(bench.PKG); D = importlib.import_module(bench.PKG + ".data")
dev = torch.device("cuda", 0)
r = bench.hbm_kernel_rooflines(pkg, torch, dev, D.default_model_config())
for x in r: print(f"{x['frac']:.3f} {x['us_per_launch']:8.1f} us  {x['kernel'][:90]}")
json.dump(r, open('gpurun_out/r2_hbm_kernels.json', 'w'), indent=1)
PY
