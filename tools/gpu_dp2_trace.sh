set -x
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29721"
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_dp.py -x -q -m gpu > gpurun_out/pytest_dp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dp.log; tail -4 gpurun_out/pytest_dp.log
timeout 300 $TR tools/dp_trace.py peer 16 gpurun_out/trace_peer_graph 2>&1 | grep "^\[rank"
M2F_EXCHANGE_IN_GRAPH=0 timeout 300 $TR tools/dp_trace.py peer 16 gpurun_out/trace_peer_cb 2>&1 | grep "^\[rank"
show() { python - "$1" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l)
        print(sys.argv[1], 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), 'e2e', d.get('e2e') and round(d['e2e']['ms_per_step'], 3), d['config'].get('exchange_launch'), d['config'].get('exchange_calls_per_step'))
PY
}
run() { name=$1; shift; timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench${N}_$name.json 2> gpurun_out/bench${N}_$name.err; show gpurun_out/bench${N}_$name.json; }
run peer_graph --exchange peer
run peer_graph8 --exchange peer --no-e2e --bucket-mb 8
M2F_EXCHANGE_IN_GRAPH=0 run peer_cb --exchange peer --no-e2e
