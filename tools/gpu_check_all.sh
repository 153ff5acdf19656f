# usage: bash tools/gpu_check_all.sh [ngpus]   -- GPU test suite, 1-GPU bench, N-GPU bench (peer and NCCL exchange)
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
show() { python - "$1" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l)
        print(sys.argv[1], 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), 'e2e', d.get('e2e') and round(d['e2e']['ms_per_step'], 3), 'roof', d.get('roofline') and round(d['roofline']['frac'], 3))
PY
}
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; show gpurun_out/bench1.json
if [ "$N" -gt 1 ]; then
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29731"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --exchange peer > gpurun_out/bench${N}_peer.json 2> gpurun_out/bench${N}_peer.err; show gpurun_out/bench${N}_peer.json
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --exchange nccl > gpurun_out/bench${N}_nccl.json 2> gpurun_out/bench${N}_nccl.err; show gpurun_out/bench${N}_nccl.json
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --exchange peer --bucket-mb 32 > gpurun_out/bench${N}_peer32.json 2> gpurun_out/bench${N}_peer32.err; show gpurun_out/bench${N}_peer32.json
fi
