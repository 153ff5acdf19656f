"""One or two launches of every kernel class the rooflines quote, at the sizes bench.py times them -- the command ncu profiles
(tools/gpu_r2_ncu.sh):  python tools/kernel_loop.py [reps]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
ops = pkg.ops
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
M, N, K = 2112, 2048, 768

# GEMM classes of FFN1 (bench.py gemm_roofline): forward with 2 planes (3 plane products), dgrad and wgrad with 1
A2 = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), 2)
W2 = ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, 2)
Y2 = ops.PlanesTensor.empty(M, N, 2, dev)
bias = torch.zeros(N, device=dev)
dH = ops.PlanesTensor.from_float(torch.randn(M, N, generator=g, device=dev), 1)
W1 = ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, 1)
X1 = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), 1)
skip = torch.randn(M, K, generator=g, device=dev)
dX = torch.empty(M, K, device=dev)
dW = torch.empty(N, K, device=dev)
for _ in range(reps):
    ops.gemm(A2, W2, M, N, bias=bias, relu=True, drop_p=0.4, seed=1, step=1, site=7, out_planes=Y2)
    ops.gemm(dH, W1, M, K, b_mn=True, resid=skip, out_f32=dX)
    ops.gemm(dH, X1, N, K, a_mn=True, b_mn=True, out_f32=dW)
torch.cuda.synchronize()

# 256 x 192 pair tiles (the plan's choice for the 768- / 2304-wide forward GEMMs): out-projection forward, 2 planes, dropout + residual
import ctypes as C
Lh = pkg._lib.lib()
Lh.m2f_debug_force_gemm_bn.argtypes = [C.c_int]
Lh.m2f_debug_force_gemm_bn.restype = None
Ao = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device=dev), 2)
Wo = ops.PlanesTensor.from_float(torch.randn(K, K, generator=g, device=dev) * K ** -0.5, 2)
bo, ro, yo = torch.zeros(K, device=dev), torch.randn(M, K, generator=g, device=dev), torch.empty(M, K, device=dev)
Lh.m2f_debug_force_gemm_bn(192)
for _ in range(reps):
    ops.gemm(Ao, Wo, M, K, bias=bo, resid=ro, out_f32=yo, drop_p=0.4, seed=1, step=1, site=8)
Lh.m2f_debug_force_gemm_bn(-1)
torch.cuda.synchronize()

# the persistent two-accumulator GEMM at the C5 inference size (FFN1, one plane)
Mi = 131072
Ai = ops.PlanesTensor.from_float(torch.randn(Mi, K, generator=g, device=dev), 1)
Yi = ops.PlanesTensor.empty(Mi, N, 1, dev)
for _ in range(reps):
    ops.gemm(Ai, W1, Mi, N, bias=bias, relu=True, out_planes=Yi)
torch.cuda.synchronize()
del Ai, Yi

# LayerNorm, staged kernels (M = 131072)
Ml, D = 131072, 768
x = torch.randn(Ml, D, generator=g, device=dev)
sk = torch.randn(Ml, D, generator=g, device=dev)
gamma, beta = torch.ones(D, device=dev), torch.zeros(D, device=dev)
y = torch.empty(Ml, D, device=dev)
ypl = ops.PlanesTensor.empty(Ml, D, 1, dev)
stats = torch.empty(Ml, 2, device=dev)
dgamma, dbeta, dbias = (torch.zeros(D, device=dev) for _ in range(3))
for _ in range(reps):
    ops.ln_fwd(x, gamma, beta, skip=sk, y=y, y_planes=ypl, drop_p=0.4, seed=1, step=2, site=3, stats=stats)
    ops.ln_bwd(sk, x, gamma, stats, dx=y, dx_planes=ypl, drop_p=0.4, seed=4, step=5, site=6, dgamma=dgamma, dbeta=dbeta, dbias=dbias)
torch.cuda.synchronize()
del x, sk, y, ypl

# LayerNorm at the training size (M = 2112): the register-resident kernels
xs, gs = torch.randn(M, D, generator=g, device=dev), torch.randn(M, D, generator=g, device=dev)
ys, yps, sts = torch.empty(M, D, device=dev), ops.PlanesTensor.empty(M, D, 1, dev), torch.empty(M, 2, device=dev)
for _ in range(reps):
    ops.ln_fwd(xs, gamma, beta, y=ys, y_planes=yps, stats=sts)
    ops.ln_bwd(gs, xs, gamma, sts, dx=ys, dx_planes=yps, drop_p=0.4, seed=4, step=5, site=6, dgamma=dgamma, dbeta=dbeta, dbias=dbias)
torch.cuda.synchronize()

# attention: training kernels (S = 33, 2 planes forward / 1 plane backward) and the long-dialogue inference kernel
Bq, Sq, H = 2048, 33, 8
Mq = Bq * Sq
mk = lambda P: ops.PlanesTensor.from_float(torch.randn(Mq, D, generator=g, device=dev) * 0.5, P)
q2, k2, v2 = mk(2), mk(2), mk(2)
o2 = ops.PlanesTensor.empty(Mq, D, 2, dev)
q1, k1, v1, do1 = mk(1), mk(1), mk(1), mk(1)
dq, dk, dv = (ops.PlanesTensor.empty(Mq, D, 1, dev) for _ in range(3))
pad = torch.zeros(Bq, Sq, dtype=torch.uint8, device=dev)
for _ in range(reps):
    ops.attn_fwd(q2, k2, v2, pad, Bq, Sq, H, o2, drop_p=0.4, seed=1, step=1, site=5)
    ops.attn_bwd(q1, k1, v1, do1, pad, Bq, Sq, H, dq, dk, dv, drop_p=0.4, seed=1, step=1, site=5)
torch.cuda.synchronize()
del q2, k2, v2, o2, q1, k1, v1, do1, dq, dk, dv
Bl, Sl = 512, 256
Mlg = Bl * Sl
mk = lambda: ops.PlanesTensor.from_float(torch.randn(Mlg, D, generator=g, device=dev) * 0.5, 1)
q, k, v = mk(), mk(), mk()
o = ops.PlanesTensor.empty(Mlg, D, 1, dev)
padl = torch.zeros(Bl, Sl, dtype=torch.uint8, device=dev)
for _ in range(reps):
    ops.attn_fwd(q, k, v, padl, Bl, Sl, H, o)
torch.cuda.synchronize()
print("kernel_loop done")
