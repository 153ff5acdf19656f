"""Isolated GEMM launches for profiling: python tools/gemm_loop.py M N K [planes] [iters] [mode: tn|ffn1|dgrad|wgrad]
(ffn1 = tn with the FFN1 epilogue: bias + ReLU + planes output, the shape class bench.py's roofline uses)"""
import importlib
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
ops = pkg.ops
M, N, K = (int(x) for x in sys.argv[1:4])
planes = int(sys.argv[4]) if len(sys.argv) > 4 else 1
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 20
mode = sys.argv[6] if len(sys.argv) > 6 else "tn"
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(1)
if mode in ("tn", "ffn1"):
    A = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), planes)
    Bm = ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev), planes)
    kw = {}
elif mode == "dgrad":  # D[M,N] = A[M,K] . B[K,N]
    A = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), planes)
    Bm = ops.PlanesTensor.from_float(torch.randn(K, N, generator=g, device=dev), planes)
    kw = dict(b_mn=True)
else:  # wgrad: D[M,N] = A[K,M]^T . B[K,N]
    A = ops.PlanesTensor.from_float(torch.randn(K, M, generator=g, device=dev), planes)
    Bm = ops.PlanesTensor.from_float(torch.randn(K, N, generator=g, device=dev), planes)
    kw = dict(a_mn=True, b_mn=True)
out = torch.empty(M, N, device=dev)
if mode == "ffn1":
    kw = dict(bias=torch.zeros(N, device=dev), relu=True, out_planes=ops.PlanesTensor.empty(M, N, planes, dev))
else:
    kw["out_f32"] = out
for _ in range(3):
    ops.gemm(A, Bm, M, N, **kw)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(iters):
    ops.gemm(A, Bm, M, N, **kw)
e1.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / iters
print(f"{mode} {M}x{N}x{K} planes {planes}: {us:.1f} us/launch (host enqueue {1e6*(t1-t0)/iters:.1f} us), {2*M*N*K/us/1e6:.1f} TFLOP/s")
