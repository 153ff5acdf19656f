"""Phase timeline of the tcgen05 long-dialogue attention (CTA 0, its second item; %globaltimer): python tools/atc_trace.py"""
import ctypes
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
Bl, Sl, H, Dm = 512, 256, 8, 768
Ml = Bl * Sl
mk = lambda: ops.PlanesTensor.from_float(torch.randn(Ml, Dm, generator=g, device=dev) * 0.5, 1)
q, k, v = mk(), mk(), mk()
o = ops.PlanesTensor.empty(Ml, Dm, 1, dev)
pad = torch.zeros(Bl, Sl, dtype=torch.uint8, device=dev)
for _ in range(3):
    ops.attn_fwd(q, k, v, pad, Bl, Sl, H, o)
torch.cuda.synchronize()
tr = torch.zeros(32, dtype=torch.int64, device=dev)
L = ctypes.CDLL(pkg._lib.LIB_PATH)
L.m2f_debug_set_attn_trace.argtypes = [ctypes.c_void_p]
L.m2f_debug_set_attn_trace(tr.data_ptr())
ops.attn_fwd(q, k, v, pad, Bl, Sl, H, o)
torch.cuda.synchronize()
L.m2f_debug_set_attn_trace(None)
t = tr.cpu().tolist()
names = {0: "mma: item start", 1: "mma: Q0,K landed", 2: "mma: TMEM tile 0 free", 3: "mma: QK0 issued", 4: "mma: Q1 landed", 5: "mma: TMEM tile 1 free",
         6: "mma: QK1 issued", 7: "mma: scores complete (prefetch K', Q0')", 8: "mma: P0 ready", 9: "mma: PV0 issued", 10: "mma: P1 ready",
         11: "mma: PV1 issued", 12: "mma: O1 complete (prefetch V')", 16: "sm: item start (key mask done)", 17: "sm: S0 complete",
         18: "sm: tile0 max exchanged", 19: "sm: tile0 may write P0", 20: "sm: tile0 P0 written", 21: "sm: tile0 past P0 barrier",
         22: "sm: S1 complete", 23: "sm: tile1 max exchanged", 24: "sm: tile1 may write P1 (O0 complete)", 25: "sm: tile1 P1 written",
         26: "sm: tile1 past P1 barrier", 28: "out: O0 complete", 29: "out: O0 stored", 30: "out: O1 complete", 31: "out: O1 stored"}
t0 = min(x for x in t if x > 0)
for i, x in sorted(((i, x) for i, x in enumerate(t) if x > 0), key=lambda kv: kv[1]):
    print(f"  +{(x - t0) / 1e3:7.2f} us  [{i:2d}] {names.get(i, '?')}")
