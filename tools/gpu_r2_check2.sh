#!/bin/bash
# round 2: all GPU tests, the bandwidth-bound kernels' rooflines, the default bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1200 -p no:cacheprovider -rA 2>&1 | grep -v "^$" | tail -250 > gpurun_out/r2_pytest_gpu.log
tail -4 gpurun_out/r2_pytest_gpu.log
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
python bench.py > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err
echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2_bench_1gpu.json').read().strip().splitlines()[-1])
print('ms', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], d['roofline']['us_per_launch'])
for g in d['roofline']['gemm_classes']: print('  ', round(g['frac'], 3), round(g['us_per_launch'], 2), g['kernel'][:70])
print({k: round(v['ms_per_step'], 3) for k, v in d['modes'].items()})
print('c5', d['extra_workloads']['c5_infer']['ms_per_batch'], 'c4', d['extra_workloads']['c4_epoch']['valid_utterances_per_s'])
PY
