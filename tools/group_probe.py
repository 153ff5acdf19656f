"""Would one persistent launch over the tiles of BOTH modality stacks beat the two concurrent launches of today?  Emulation with what
exists: (a) the same GEMM on two streams at M = 2112 (today), (b) one launch at M = 4224 (M2F_GEMM_PERSIST=2 forces the persistent
kernel at any tile count).  Usage: python tools/group_probe.py  ->  one line per shape, us per pair of GEMMs."""
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
M = 2112
REPS = 40
shapes = [("qkv", 2304, 768, "pl"), ("wo", 768, 768, "res"), ("ffn1", 2048, 768, "relu"), ("ffn2", 768, 2048, "res")]
s0, s1 = torch.cuda.Stream(), torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for planes in (2, 1):
    for name, N, K, kind in shapes:
        def operands(rows):
            A = ops.PlanesTensor.from_float(torch.randn(rows, K, generator=g, device=dev), planes)
            W = ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device=dev) * K ** -0.5, planes)
            b = torch.zeros(N, device=dev)
            kw = dict(bias=b)
            if kind == "res":
                kw.update(resid=torch.randn(rows, N, generator=g, device=dev), out_f32=torch.empty(rows, N, device=dev))
            else:
                kw.update(out_planes=ops.PlanesTensor.empty(rows, N, planes, dev), relu=(kind == "relu"))
            return A, W, kw
        a0, a1, big = operands(M), operands(M), operands(2 * M)

        def two():
            with torch.cuda.stream(s0):
                ops.gemm(a0[0], a0[1], M, N, **a0[2])
            with torch.cuda.stream(s1):
                ops.gemm(a1[0], a1[1], M, N, **a1[2])

        def one():
            with torch.cuda.stream(s0):
                ops.gemm(big[0], big[1], 2 * M, N, **big[2])

        res = {}
        for label, fn in (("two_streams_M2112", two), ("one_launch_M4224", one)):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            cap = torch.cuda.Stream()
            with torch.cuda.graph(graph, stream=cap):   # a chain of REPS fork / join rounds, replayed as one graph (no host time inside)
                for _ in range(REPS):
                    s0.wait_stream(cap)
                    s1.wait_stream(cap)
                    fn()
                    cap.wait_stream(s0)
                    cap.wait_stream(s1)
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
            res[label] = e0.elapsed_time(e1) / (5 * REPS) * 1e3
        print(f"planes {planes} {name:5s} N {N} K {K}: " + "  ".join(f"{k} {v:7.1f} us" for k, v in res.items()), flush=True)
