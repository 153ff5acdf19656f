# usage: bash tools/gpu_benchN.sh N   -- N-GPU bench, peer exchange (full line) then NCCL exchange (no e2e)
set -x
N=${1:-8}
mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l)
        print(sys.argv[1], 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), 'e2e', d.get('e2e') and round(d['e2e']['ms_per_step'], 3), d['config'].get('exchange'), d['config'].get('exchange_calls_per_step'))
PY
}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741"
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --exchange peer > gpurun_out/bench${N}_peer.json 2> gpurun_out/bench${N}_peer.err; show gpurun_out/bench${N}_peer.json; tail -3 gpurun_out/bench${N}_peer.err
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --exchange nccl > gpurun_out/bench${N}_nccl.json 2> gpurun_out/bench${N}_nccl.err; show gpurun_out/bench${N}_nccl.json
