"""Minimal training-step loop for profiling: python tools/step_loop.py [steps] [precision] [batch] [seq]
(same step as bench.py: fwd + fused CE + bwd + fused Adam, dropout 0.4, default config)."""
import importlib
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
D = importlib.import_module("multimodal-emotion-recognition_b200.data")

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
precision = sys.argv[2] if len(sys.argv) > 2 else "bf16x2f"
B = int(sys.argv[3]) if len(sys.argv) > 3 else 64
S = int(sys.argv[4]) if len(sys.argv) > 4 else 33
torch.manual_seed(0)
model = pkg.M2FNet(D.default_model_config(), precision=precision).to("cuda").train()
crit = pkg.FusedCrossEntropyLoss().cuda()
opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
b = {k: v.cuda() for k, v in D.synth_batch(B, S, 768, 768).items()}
L = pkg._lib.lib()
for i in range(steps):
    torch.cuda.synchronize()
    t0, l0 = time.perf_counter(), L.m2f_launch_count()
    opt.zero_grad(set_to_none=True)
    loss = crit(model(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"])
    t1 = time.perf_counter()
    loss.backward()
    t2 = time.perf_counter()
    opt.step()
    t3 = time.perf_counter()
    torch.cuda.synchronize()
    t4 = time.perf_counter()
    print(f"step {i}: host enqueue fwd {1e3*(t1-t0):.2f} ms, bwd {1e3*(t2-t1):.2f} ms, adam {1e3*(t3-t2):.2f} ms; total with sync {1e3*(t4-t0):.2f} ms; "
          f"launches {L.m2f_launch_count()-l0}; loss {loss.item():.4f}", flush=True)
