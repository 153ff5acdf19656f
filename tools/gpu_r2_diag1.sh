#!/bin/bash
# round 2: static-scratch CE test (alone, then after the op tests) + per-tensor diagnosis of the train-mode parity case
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/r2_zz_alone.log
python -m pytest tests/test_gpu_ops.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/r2_zz_after.log
tail -3 gpurun_out/r2_zz_alone.log gpurun_out/r2_zz_after.log
python tools/diag_dropout.py > gpurun_out/r2_diag_default.log 2>&1
M2F_NARROW=0 python tools/diag_dropout.py > gpurun_out/r2_diag_narrow0.log 2>&1
M2F_LN_V2=0 python tools/diag_dropout.py > gpurun_out/r2_diag_ln0.log 2>&1
M2F_GRAPH=0 M2F_STREAMS=0 M2F_PDL=0 python tools/diag_dropout.py > gpurun_out/r2_diag_serial.log 2>&1
M2F_GEMM_CTA2=0 python tools/diag_dropout.py > gpurun_out/r2_diag_nopairs.log 2>&1
head -30 gpurun_out/r2_diag_default.log
