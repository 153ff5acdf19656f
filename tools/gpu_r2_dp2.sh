#!/bin/bash
# round 2, 2 GPUs: the data-parallel tests the 1-GPU test box skips, then the bench line at N = 2 (exchange self-check,
# c4_epoch under DP) -- the exact launch the driver uses.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_dp2_gpus.txt
timeout 900 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -m gpu -q --timeout 600 -p no:cacheprovider -rA 2>&1 | grep -v "^$" | tail -40 > gpurun_out/r2_dp2_pytest.log
tail -12 gpurun_out/r2_dp2_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_2gpu.json 2> gpurun_out/r2_bench_2gpu.err
echo "bench2 rc=$?"
tail -c 1500 gpurun_out/r2_bench_2gpu.err
python - <<'PY'
import json
for l in open('gpurun_out/r2_bench_2gpu.json'):
    if l.startswith('{'):
        d = json.loads(l)
        print('N=2 ms', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'] and d['e2e']['value'])
        print(json.dumps({k: v for k, v in d['config'].items() if 'exchange' in k or 'sync' in k}))
        print(json.dumps(d.get('extra_workloads')))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus 2 --steps 10 --warmup 3 --impl reference > gpurun_out/r2_bench_2gpu_reference.json 2> gpurun_out/r2_bench_2gpu_reference.err
echo "ref rc=$?"; head -c 1200 gpurun_out/r2_bench_2gpu_reference.json
