#!/bin/bash
# round 2: the tcgen05 long-dialogue attention -- parity tests, then the bandwidth table (attention rows) and the C5 inference loop
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ops.py -m gpu -q -k "long_dialogues or attention" -p no:cacheprovider 2>&1 | tail -30 > gpurun_out/r2_atc_pytest.log
tail -12 gpurun_out/r2_atc_pytest.log
timeout 300 python - <<'PY'
import importlib, json, sys, torch
sys.path.insert(0, '.')
import bench
pkg = importlib.import_module(bench.PKG); D = importlib.import_module(bench.PKG + ".data")
dev = torch.device("cuda", 0)
r = bench.hbm_kernel_rooflines(pkg, torch, dev, D.default_model_config())
for x in r: print(f"{x['frac']:.3f} {x['us_per_launch']:8.1f} us  {x['kernel'][:90]}")
PY
M2F_ATT_TC=0 timeout 300 python tools/infer_loop.py 2>&1 | tail -3
timeout 300 python tools/infer_loop.py 2>&1 | tail -3
