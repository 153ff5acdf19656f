"""In-kernel timeline of one GEMM launch (globaltimer stamps per CTA): python tools/gemm_trace.py M N K [mode]"""
import ctypes
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
ops = pkg.ops
M, N, K = (int(x) for x in sys.argv[1:4])
mode = sys.argv[4] if len(sys.argv) > 4 else "tn"
g = torch.Generator(device="cuda").manual_seed(1)
kw = {}
if mode in ("tn", "ffn1"):
    A = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device="cuda"), 1)
    Bm = ops.PlanesTensor.from_float(torch.randn(N, K, generator=g, device="cuda"), 1)
elif mode == "dgrad":
    A = ops.PlanesTensor.from_float(torch.randn(M, K, generator=g, device="cuda"), 1)
    Bm = ops.PlanesTensor.from_float(torch.randn(K, N, generator=g, device="cuda"), 1)
    kw = dict(b_mn=True)
else:
    A = ops.PlanesTensor.from_float(torch.randn(K, M, generator=g, device="cuda"), 1)
    Bm = ops.PlanesTensor.from_float(torch.randn(K, N, generator=g, device="cuda"), 1)
    kw = dict(a_mn=True, b_mn=True)
out = torch.empty(M, N, device="cuda")
okw = dict(out_f32=out)
if mode == "ffn1":  # bias + ReLU + bf16 planes output (the FFN1 epilogue)
    okw = dict(bias=torch.zeros(N, device="cuda"), relu=True, out_planes=ops.PlanesTensor.empty(M, N, 1, "cuda"))
for _ in range(5):
    ops.gemm(A, Bm, M, N, **okw, **kw)
torch.cuda.synchronize()
tr = torch.zeros(2 * 4096 * 8, dtype=torch.int64, device="cuda")
L = ctypes.CDLL(pkg._lib.LIB_PATH)
L.m2f_debug_set_gemm_trace.argtypes = [ctypes.c_void_p]
L.m2f_debug_set_gemm_flags(int(os.environ.get("M2F_GEMM_DBG", "0")))
L.m2f_debug_set_gemm_trace(tr.data_ptr())
ops.gemm(A, Bm, M, N, **okw, **kw)
torch.cuda.synchronize()
L.m2f_debug_set_gemm_trace(None)
t2 = tr.cpu()[4096 * 8:].reshape(-1, 8)
t = tr.cpu()[:4096 * 8].reshape(-1, 8)
t2 = t2[t[:, 0] > 0]
t = t[t[:, 0] > 0]
t0 = t[:, 0].min()
names = ["start", "setup_done", "first_full", "mma_issued", "accum_ready", "epi_done", "all_done"]
print(f"{mode} {M}x{N}x{K}: {len(t)} CTAs; kernel span {(t[:, 6].max() - t0).item() / 1e3:.2f} us")
rel = (t[:, :7] - t[:, :1]).double()
for i, n in enumerate(names):
    print(f"  {n:12s} median +{rel[:, i].median().item() / 1e3:7.2f} us   max +{rel[:, i].max().item() / 1e3:7.2f} us")
st = (t[:, 0] - t0).double() / 1e3
print(f"  CTA start times: median {st.median().item():.2f} us, max {st.max().item():.2f} us")
print("  epilogue chunk 0 cycles (tmem_ld, stage, process):", [int(t2[:, i].double().median().item()) for i in range(3)],
      " chunk 1:", [int(t2[:, 4 + i].double().median().item()) for i in range(3)], " chunk0->chunk1 start:", int((t2[:, 7] - t2[:, 3]).double().median().item()))
