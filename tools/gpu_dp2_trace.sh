set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721"
timeout 300 $TR tools/dp_trace.py peer 32 gpurun_out/trace_peer 2>&1 | grep -v "^\*\|OMP_NUM" | tail -8
timeout 300 $TR tools/dp_trace.py nccl 128 gpurun_out/trace_nccl 2>&1 | grep -v "^\*\|OMP_NUM" | tail -8
run() { name=$1; shift; timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e "$@" > gpurun_out/bench2_$name.json 2> gpurun_out/bench2_$name.err; echo "$name rc=$?"; python -c "
import json,sys
for l in open('gpurun_out/bench2_$name.json'):
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],3))
"; }
M2F_EXCHANGE_CTAS=8 run peer_c8 --exchange peer
M2F_EXCHANGE_CTAS=16 run peer_c16 --exchange peer
M2F_EXCHANGE_CTAS=64 run peer_c64 --exchange peer
M2F_EXCHANGE_DRYRUN=1 run peer_dry --exchange peer
M2F_EXCHANGE_MC=0 run peer_p2p --exchange peer
run peer_b64 --exchange peer --bucket-mb 64
