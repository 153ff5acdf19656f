"""Inference throughput of the test() body (src/test.py:58-71: eval, inference_mode, argmax): python tools/infer_loop.py [B] [S] [precision]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
D = importlib.import_module("multimodal-emotion-recognition_b200.data")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
S = int(sys.argv[2]) if len(sys.argv) > 2 else 256
planes = 3 if (len(sys.argv) > 3 and sys.argv[3] == "fp32") else 1
torch.manual_seed(0)
model = pkg.M2FNet(D.default_model_config(), planes=planes).to("cuda").eval()
b = {k: v.cuda() for k, v in D.synth_batch(B, S, 768, 768).items()}
L = pkg._lib.lib()
with torch.inference_mode():
    for _ in range(2):
        pred = model(b["text"], b["audio"], b["padding_mask"]).argmax(dim=2)
    torch.cuda.synchronize()
    print("workspace MB:", sum(p.ws_bytes for p in model._plans.values()) / 1e6, flush=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    l0 = L.m2f_launch_count()
    e0.record()
    for _ in range(n):
        pred = model(b["text"], b["audio"], b["padding_mask"]).argmax(dim=2)
    e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
flop = 185.6e6 * B * S if S == 256 else 174.0e6 * B * S
print(f"B={B} S={S} planes={planes}: {ms:.2f} ms/batch, {B*S/ms*1e3:.0f} utt/s, {flop/ms/1e9:.1f} TFLOP/s, launches/batch {(L.m2f_launch_count()-l0)//n}")
