#!/bin/bash
# round 2, GPU call 1: every -m gpu test (new: train-mode dropout parity, SURVEY 8c protocol, bf16x2f mode, staged LayerNorm,
# unchanged reference drivers) and the default bench line.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/r2_gpu_info.txt 2>&1
python -m pytest tests -m gpu -q --timeout 1200 -p no:cacheprovider -rA 2>&1 | grep -v "^$" | tail -250 > gpurun_out/r2_pytest_gpu.log
tail -5 gpurun_out/r2_pytest_gpu.log
python bench.py > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err
echo "bench rc=$?"
tail -c 3000 gpurun_out/r2_bench_1gpu.err
head -c 6000 gpurun_out/r2_bench_1gpu.json
