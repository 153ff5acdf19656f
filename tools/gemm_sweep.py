"""GEMM timings for the model's shapes with launch overhead removed (CUDA-graph replay of 20 launches, operands
rotated over 6 sets so consecutive launches read different weights): python tools/gemm_sweep.py [planes] [M]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
ops = pkg.ops
planes = int(sys.argv[1]) if len(sys.argv) > 1 else 1
M = int(sys.argv[2]) if len(sys.argv) > 2 else 2112
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(1)
SETS, REP = 6, 20


def rnd(r, c):
    return ops.PlanesTensor.from_float(torch.randn(r, c, generator=g, device=dev), planes)


def bench(name, mode, m, n, k, out_planes=False):
    if mode == "tn":
        A = [rnd(m, k) for _ in range(SETS)]
        B = [rnd(n, k) for _ in range(SETS)]
        kw = {}
    elif mode == "dgrad":
        A = [rnd(m, k) for _ in range(SETS)]
        B = [rnd(k, n) for _ in range(SETS)]
        kw = dict(b_mn=True)
    else:
        A = [rnd(k, m) for _ in range(SETS)]
        B = [rnd(k, n) for _ in range(SETS)]
        kw = dict(a_mn=True, b_mn=True)
    if out_planes:
        outs = [ops.PlanesTensor.empty(m, n, planes, dev) for _ in range(SETS)]
        okw = [dict(out_planes=o) for o in outs]
    else:
        outs = [torch.empty(m, n, device=dev) for _ in range(SETS)]
        okw = [dict(out_f32=o) for o in outs]
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(SETS):
            ops.gemm(A[i], B[i], m, n, **kw, **okw[i])
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            for i in range(REP):
                ops.gemm(A[i % SETS], B[i % SETS], m, n, **kw, **okw[i % SETS])
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(5):
            gr.replay()
        e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (5 * REP)
    print(f"{name:28s} {mode:6s} {m:6d}x{n:5d}x{k:5d} {'planes' if out_planes else 'f32   '} {us:7.2f} us  {2.0 * m * n * k / us / 1e6:7.1f} TFLOP/s", flush=True)
    return us


tot = 0.0
tot += 12 * bench("enc qkv fwd", "tn", M, 2304, 768, True)
tot += 24 * bench("enc out-proj / fam fwd", "tn", M, 768, 768, False)
tot += 12 * bench("enc ffn1 fwd", "tn", M, 2048, 768, True)
tot += 12 * bench("enc ffn2 fwd", "tn", M, 768, 2048, False)
tot += 12 * bench("dgrad ffn2 (dh)", "dgrad", M, 2048, 768, True)
tot += 12 * bench("dgrad ffn1", "dgrad", M, 768, 2048, False)
tot += 30 * bench("dgrad 768x768", "dgrad", M, 768, 768, True)
tot += 12 * bench("dgrad qkv", "dgrad", M, 768, 2304, False)
tot += 34 * bench("wgrad 768x768", "wgrad", 768, 768, M)
tot += 12 * bench("wgrad qkv", "wgrad", 2304, 768, M)
tot += 12 * bench("wgrad ffn1", "wgrad", 2048, 768, M)
tot += 12 * bench("wgrad ffn2", "wgrad", 768, 2048, M)
print(f"weighted total (approximate step GEMM time): {tot / 1e3:.2f} ms")
