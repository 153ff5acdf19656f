"""Per-launch time of the small kernels of the step at the training size (M = 2112), measured as a chain of REPS dependent launches
replayed as one CUDA graph (no host time inside): what a kernel costs on the critical path when nothing else runs beside it.
Usage: python tools/chain_probe.py"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
ops = pkg.ops
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
M, D, B, S, H = 2112, 768, 64, 33, 8
REPS = 40

x, sk = torch.randn(M, D, generator=g, device=dev), torch.randn(M, D, generator=g, device=dev)
gamma, beta = torch.ones(D, device=dev), torch.zeros(D, device=dev)
y, stats = torch.empty(M, D, device=dev), torch.empty(M, 2, device=dev)
dgamma, dbeta, dbias = (torch.zeros(D, device=dev) for _ in range(3))
pad = torch.zeros(B, S, dtype=torch.uint8, device=dev)
mk = lambda P: ops.PlanesTensor.from_float(torch.randn(M, D, generator=g, device=dev) * 0.5, P)
cases = {}
for P in (2, 1):
    ypl = ops.PlanesTensor.empty(M, D, P, dev)
    cases[f"ln_fwd planes {P}"] = lambda ypl=ypl: ops.ln_fwd(x, gamma, beta, y=y, y_planes=ypl, drop_p=0.4, seed=1, step=2, site=3, stats=stats)
ypl1 = ops.PlanesTensor.empty(M, D, 1, dev)
ops.ln_fwd(x, gamma, beta, y=y, y_planes=ypl1, stats=stats)
cases["ln_bwd planes 1 (dropout, dbias)"] = lambda: ops.ln_bwd(sk, x, gamma, stats, dx=y, dx_planes=ypl1, drop_p=0.4, seed=4, step=5, site=6,
                                                                dgamma=dgamma, dbeta=dbeta, dbias=dbias)
cases["ln_bwd no planes (final norm)"] = lambda: ops.ln_bwd(sk, x, gamma, stats, dx=y, dgamma=dgamma, dbeta=dbeta)
q2, k2, v2, o2 = mk(2), mk(2), mk(2), ops.PlanesTensor.empty(M, D, 2, dev)
cases["attn_fwd planes 2 (dropout)"] = lambda: ops.attn_fwd(q2, k2, v2, pad, B, S, H, o2, drop_p=0.4, seed=1, step=1, site=5)
q1, k1, v1, do1 = mk(1), mk(1), mk(1), mk(1)
dq, dk, dv = (ops.PlanesTensor.empty(M, D, 1, dev) for _ in range(3))
dbq = torch.zeros(3 * D, device=dev)
cases["attn_bwd planes 1 (dropout, bias sums)"] = lambda: ops.attn_bwd(q1, k1, v1, do1, pad, B, S, H, dq, dk, dv, drop_p=0.4, seed=1, step=1, site=5,
                                                                         dbq=dbq[:D], dbk=dbq[D:2 * D], dbv=dbq[2 * D:])
sp_in = torch.randn(M, D, generator=g, device=dev)
sp_out = ops.PlanesTensor.empty(M, D, 2, dev)
cases["split_planes 2"] = lambda: pkg._lib.check(pkg._lib.lib().m2f_split_planes(sp_in.data_ptr(), D, __import__("ctypes").byref(sp_out.c()),
                                                                                  torch.cuda.current_stream().cuda_stream), "split")
for name, fn in cases.items():
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(REPS):
            fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:45s} {e0.elapsed_time(e1) / (5 * REPS) * 1e3:7.2f} us per launch in a chain", flush=True)
