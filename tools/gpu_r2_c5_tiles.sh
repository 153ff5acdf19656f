#!/bin/bash
# round 2: C5 inference batch under the three GEMM tile policies
for v in 1 0 2; do echo "M2F_GEMM_CTA2=$v"; M2F_GEMM_CTA2=$v timeout 300 python tools/infer_loop.py 2>&1 | tail -1; done
echo "pair96-like: M2F_NARROW=0"; M2F_NARROW=0 timeout 300 python tools/infer_loop.py 2>&1 | tail -1
