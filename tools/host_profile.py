"""cProfile of the host side of the training step (what the Python layer costs per step): python tools/host_profile.py"""
import cProfile
import importlib
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
D = importlib.import_module("multimodal-emotion-recognition_b200.data")
torch.manual_seed(0)
model = pkg.M2FNet(D.default_model_config()).to("cuda").train()
crit = pkg.FusedCrossEntropyLoss().cuda()
opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
b = {k: v.cuda() for k, v in D.synth_batch(64, 33, 768, 768).items()}


def step():
    opt.zero_grad(set_to_none=True)
    loss = crit(model(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"])
    loss.backward()
    opt.step()
    return loss


for _ in range(5):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(28)
