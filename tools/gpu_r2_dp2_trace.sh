#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721"
timeout 300 $TR tools/dp_trace.py peer 16 gpurun_out/r2_trace_fp32 2>&1 | grep "^\[rank"
M2F_TRACE_WIRE=bf16 timeout 300 $TR tools/dp_trace.py peer 16 gpurun_out/r2_trace_bf16 2>&1 | grep "^\[rank"
grep -h "xchg\|end$\|begin" gpurun_out/r2_trace_fp32/dp_timeline_rank0.tsv | awk '$1>0' | head -80
echo ======= bf16
grep -h "xchg\|end$" gpurun_out/r2_trace_bf16/dp_timeline_rank0.tsv | awk '$1>0' | head -80
