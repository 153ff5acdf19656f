#!/bin/bash
# round 2, final 1-GPU validation: what the driver runs (GPU tests with -x, smoke, both bench arms)
mkdir -p gpurun_out
python -m pytest tests/ -x -q -m gpu -p no:cacheprovider 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_bench_reference_arm.err; echo "ref rc=$?"; head -c 700 gpurun_out/r2_bench_reference_arm.json; echo
( time python bench.py > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err ) 2>&1 | grep real
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2_bench_1gpu.json').read().strip().splitlines()[-1])
print('ms', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], 'traffic', d['roofline']['traffic'], 'launches', d['gpu_launches'])
print({k: round(v['ms_per_step'], 3) for k, v in d['modes'].items()}, 'clocks', d['clocks'])
print('c5', d['extra_workloads']['c5_infer']['ms_per_batch'], 'c4', d['extra_workloads']['c4_epoch']['valid_utterances_per_s'], 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['kind'])
for x in d['roofline']['hbm_kernels']: print(f"  {x['frac']:.3f} {x['kernel'][:80]}")
PY
