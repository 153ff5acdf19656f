#!/bin/bash
# per-CTA phase times of the three GEMM classes of FFN1 and a 768-wide forward (1 plane)
python tools/gemm_trace.py 2112 768 2048 dgrad 2>&1 | tail -10
python tools/gemm_trace.py 2048 768 2112 wgrad 2>&1 | tail -10
python tools/gemm_trace.py 2112 2048 768 ffn1 2>&1 | tail -10
python tools/gemm_trace.py 2112 768 768 tn 2>&1 | tail -10
