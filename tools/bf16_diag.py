import importlib, sys, torch
sys.path.insert(0, '/root/repo')
pkg = importlib.import_module("multimodal-emotion-recognition_b200")
from tests.test_gpu_model import run_case
for planes in (1, 2):
    e_logits, e_loss, e_grad, per, same = run_case(pkg, "meld_base_2l", planes)
    print("planes", planes, "logits", e_logits, "loss", e_loss, "grads", e_grad, "pred_same", same)
    for k, v in sorted(per.items(), key=lambda kv: -kv[1])[:72]:
        print(f"   {v:.3e} {k}")
