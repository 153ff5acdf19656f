# usage: bash tools/gpu_benchN.sh N [extra bench args]  -- N-GPU bench line with the default (peer) gradient exchange
set -x
N=${1:-8}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741"
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench${N}_final.json 2> gpurun_out/bench${N}_final.err; echo "rc=$?"
python - gpurun_out/bench${N}_final.json <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l)
        print('value', round(d['value']), 'ms', round(d['ms_per_step'], 3), 'e2e', d.get('e2e') and round(d['e2e']['ms_per_step'], 3), {k: v for k, v in d['config'].items() if k.startswith(('exchange', 'bucket', 'replicas'))})
PY
tail -3 gpurun_out/bench${N}_final.err
