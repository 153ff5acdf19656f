"""Time the long-dialogue attention forward alone (C5 size: 512 dialogues x 256 utterances, 8 heads of 96): python tools/atc_time.py [reps]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
ops = pkg.ops
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
Bl, Sl, H, Dm = 512, 256, 8, 768
Ml = Bl * Sl
sets = []
for _ in range(3):  # rotate operand sets: 3 x 805 MB > L2
    mk = lambda: ops.PlanesTensor.from_float(torch.randn(Ml, Dm, generator=g, device=dev) * 0.5, 1)
    sets.append((mk(), mk(), mk(), ops.PlanesTensor.empty(Ml, Dm, 1, dev)))
pad = torch.zeros(Bl, Sl, dtype=torch.uint8, device=dev)
for s in sets:
    ops.attn_fwd(s[0], s[1], s[2], pad, Bl, Sl, H, s[3])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(reps):
    s = sets[i % 3]
    ops.attn_fwd(s[0], s[1], s[2], pad, Bl, Sl, H, s[3])
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / reps
print(f"attn_fwd long (M2F_ATT_TC={os.environ.get('M2F_ATT_TC', '1')}): {us:.1f} us per MHA, {4 * Ml * Dm * 2 / us / 1e3:.0f} GB/s")
