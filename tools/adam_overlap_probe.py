"""Would the optimizer pass hide under the backward on ONE GPU?  Upper-bound probe: the fused Adam + shadow kernel runs on DUMMY
arenas on a low-priority side stream, concurrently with the real backward (no dependencies), against the serial step.
python tools/adam_overlap_probe.py [steps]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
D = importlib.import_module("multimodal-emotion-recognition_b200.data")
L = pkg._lib
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = pkg.M2FNet(D.default_model_config(), precision="bf16x2f").to(dev).train()
opt = pkg.FusedAdam(model, lr=5e-5, weight_decay=0.01)
crit = pkg.FusedCrossEntropyLoss().to(dev)
batches = [{k: v.to(dev) for k, v in D.synth_batch(64, 33, 768, 768, seed=s).items()} for s in range(8)]


def step(b, real_step=True, side=None, dummy=None):
    opt.zero_grad(set_to_none=True)
    out = model(b["text"], b["audio"], b["padding_mask"])
    loss = crit(out.permute(0, 2, 1), b["emotion"])
    if side is not None:
        ev = torch.cuda.Event()
        ev.record()
        with torch.cuda.stream(side):
            side.wait_event(ev)
            L.check(L.lib().m2f_model_adam_step(model._handle, dummy["p"].data_ptr(), dummy["g"].data_ptr(), dummy["m"].data_ptr(), dummy["v"].data_ptr(),
                                                dummy["s"].data_ptr(), 5e-5, 0.9, 0.999, 1e-8, 0.01, 1, 1.0, side.cuda_stream), "adam")
    loss.backward()
    if side is not None:
        torch.cuda.current_stream().wait_stream(side)
    if real_step:
        opt.step()


def timed(fn, n):
    for i in range(5):
        fn(batches[i % 8])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(batches[i % 8])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


base = timed(lambda b: step(b, True), steps)
nostep = timed(lambda b: step(b, False), steps)
arena = model.flat_parameters()
dummy = {"p": arena.clone(), "g": torch.randn_like(arena) * 1e-3, "m": torch.zeros_like(arena), "v": torch.zeros_like(arena), "s": model._shadow.clone()}
lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
side = torch.cuda.Stream(priority=0)
overl = timed(lambda b: step(b, False, side, dummy), steps)
print(f"fwd + bwd + step (serial): {base:.3f} ms;  fwd + bwd only: {nostep:.3f} ms;  fwd + (bwd || Adam on dummy arenas): {overl:.3f} ms")
print(f"serial Adam costs {base - nostep:.3f} ms; overlapped it costs {overl - nostep:.3f} ms")
