"""Per-tensor report of the train-mode (dropout 0.4, replayed masks) parity case meld_base_2l: top errors, q/k/v split of the packed
in-projection gradients.  Usage: [M2F_NARROW=0 ...] python tools/diag_dropout.py [case] [planes]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
P = importlib.import_module("multimodal-emotion-recognition_b200")
from oracle import m2fnet_oracle as O  # noqa: E402
from oracle.gen_golden import CASES  # noqa: E402
import test_gpu_dropout_parity as T  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "meld_base_2l"
planes = int(sys.argv[2]) if len(sys.argv) > 2 else 3
over, B, S, ragged, weighted = CASES[name]
cfg = O.default_model_cfg(dropout=T.P_DROP, **over)
params = O.synth_params(cfg)
model = P.M2FNet(cfg, planes=planes)
model.load_state_dict({k: v.float() for k, v in params.items()}, strict=True)
model = model.to("cuda").train()
b = O.synth_inputs(B, S, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size, C=cfg.CLASSIFIER.output_size)
crit = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)
for step in range(2):
    model.zero_grad(set_to_none=True)
    out = model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda())
    loss = crit(out.permute(0, 2, 1), b["emotion"].cuda())
    loss.backward()
    torch.cuda.synchronize()
    masks = T.SiteMasks(P, model._seed, model._step)
    ref_logits, ref_loss, ref_grads = O.loss_and_grads(params, cfg, b["text"].double(), b["audio"].double(), b["padding_mask"],
                                                       b["emotion"], None, drop_masks=masks, p=T.P_DROP)
    got = {O.canonical_name(n): q.grad.detach().cpu().double() for n, q in model.named_parameters()}
    per = {n: O.rel_l2(got[n], ref_grads[n]) for n in ref_grads if float(ref_grads[n].abs().max()) > 0}
    print(f"step {step} env NARROW={os.environ.get('M2F_NARROW')} LN_V2={os.environ.get('M2F_LN_V2')}: top errors")
    for n, e in sorted(per.items(), key=lambda kv: -kv[1])[:12]:
        print(f"   {e:.3e}  {n}  |ref| {float(torch.linalg.vector_norm(ref_grads[n])):.3e}")
    for n in per:
        if "in_proj" in n and ("fusion" in n or per[n] > 1e-4):
            D = got[n].shape[0] // 3
            parts = []
            for i, tag in enumerate("qkv"):
                g, r = got[n][i * D:(i + 1) * D], ref_grads[n][i * D:(i + 1) * D]
                nr = float(torch.linalg.vector_norm(r))
                parts.append(f"{tag}: err {float(torch.linalg.vector_norm(g - r)) / max(nr, 1e-300):.2e} |ref| {nr:.2e}")
            print("   split", n, " ".join(parts))
