"""Generate tests/golden/*.npz from the UNMODIFIED reference module.

Run in the build container only (needs /root/reference, which does not exist on the GPU
box):   python oracle/gen_golden.py

For every case it imports /root/reference/src/model.py by path, constructs
`M2FNet(config)`, strict-loads RNG-free weights (`oracle.m2fnet_oracle.synth_params`, so
the fixture does not have to store them), runs forward + `CrossEntropyLoss(weight?,
ignore_index=-1, label_smoothing=0.1)` on the permuted view + backward exactly as
/root/reference/src/train.py:228-230 does, in fp64 with dropout 0 (train() == eval() bit
for bit at p = 0), and stores inputs, logits, loss and gradients.  Large
gradient tensors are stored as (sum, l2, head, strided sample) to keep fixtures small.
TEST INFRASTRUCTURE -- not imported by the product.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import m2fnet_oracle as O  # noqa: E402

REF_MODEL = "/root/reference/src/model.py"
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
FULL_LIMIT = 4096
SAMPLE = 257

CASES = {
    # name: (cfg overrides, B, S, ragged, weighted CE)
    "both_fam": (dict(AUDIO=dict(embedding_size=48, n_head=4, n_encoder_layers=2),
                      TEXT=dict(embedding_size=64, n_head=4, n_encoder_layers=2),
                      FAM=dict(embedding_size=32, n_head=2, n_layers=2),
                      CLASSIFIER=dict(hidden_size=40, output_size=7, n_layers=2)), 4, 9, True, False),
    "shared_norm_cls3_weighted": (dict(AUDIO=dict(embedding_size=60, n_head=6, n_transformers=2, n_encoder_layers=1),
                                       TEXT=dict(embedding_size=32, n_head=2, n_transformers=2, n_encoder_layers=2),
                                       FAM=dict(embedding_size=48, n_head=3, n_layers=3),
                                       CLASSIFIER=dict(hidden_size=24, output_size=7, n_layers=3)), 3, 7, True, True),
    "audio_only": (dict(AUDIO=dict(embedding_size=32, n_head=4, n_encoder_layers=1),
                        TEXT=dict(enabled=False, embedding_size=32, n_head=4, n_encoder_layers=1),
                        FAM=dict(enabled=False, embedding_size=40, n_head=4, n_layers=1),
                        CLASSIFIER=dict(hidden_size=16, output_size=7, n_layers=2)), 2, 5, True, False),
    "text_only": (dict(AUDIO=dict(enabled=False, embedding_size=32, n_head=4, n_encoder_layers=1),
                       TEXT=dict(embedding_size=40, n_head=5, n_encoder_layers=1),
                       FAM=dict(enabled=False, embedding_size=24, n_head=4, n_layers=1),
                       CLASSIFIER=dict(hidden_size=16, output_size=7, n_layers=1)), 2, 6, True, True),
    "concat_no_fam": (dict(AUDIO=dict(embedding_size=32, n_head=2, n_encoder_layers=1),
                           TEXT=dict(embedding_size=48, n_head=3, n_encoder_layers=1),
                           FAM=dict(enabled=False, embedding_size=32, n_head=4, n_layers=1),
                           CLASSIFIER=dict(hidden_size=32, output_size=5, n_layers=4)), 3, 4, False, False),
    # MELD-sized widths (text_base 768 + wav2vec2 768, 8 heads, FFN 2048), 2 + 2 encoder layers, 2 fusion layers: the case the
    # bf16 tolerance of BASELINE.json is judged on.  Inputs are RNG-free (O.synth_inputs) and not stored.
    "meld_base_2l": (dict(AUDIO=dict(n_encoder_layers=2), TEXT=dict(n_encoder_layers=2), FAM=dict(n_layers=2)), 4, 33, "rngfree", False),
}


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_model", REF_MODEL)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def case_inputs(cfg, B, S, ragged, weighted, seed):
    C = cfg.CLASSIFIER.output_size
    if ragged == "rngfree":
        b = O.synth_inputs(B, S, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size, C=C)
    else:
        b = O.synth_batch(B, S, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size, C=C, seed=seed, ragged=ragged)
    w = None
    if weighted:
        p = torch.tensor([.47, .17, .07, .11, .12, .03, .03], dtype=torch.float64)[:C]
        w = 1.0 / (C * p / p.sum())
    return b, w


def summarise(g: torch.Tensor):
    f = g.detach().double().flatten()
    n = f.numel()
    if n <= FULL_LIMIT:
        return {"full": f.numpy()}
    step = max(1, n // SAMPLE)
    return {"sum": np.array(f.sum().item()), "l2": np.array(torch.linalg.vector_norm(f).item()),
            "head": f[:32].numpy(), "sample": f[::step].numpy(), "step": np.array(step)}


def main():
    warnings.filterwarnings("ignore")
    ref = load_reference()
    os.makedirs(OUT, exist_ok=True)
    for ci, (name, (over, B, S, ragged, weighted)) in enumerate(CASES.items()):
        cfg = O.default_model_cfg(dropout=0.0, **over)
        model = ref.M2FNet(cfg).double()
        params = O.synth_params(cfg)
        missing = model.load_state_dict(params, strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys
        batch, w = case_inputs(cfg, B, S, ragged, weighted, seed=100 + ci)
        text, audio = batch["text"].double(), batch["audio"].double()
        mask, emo = batch["padding_mask"], batch["emotion"]
        model.train()
        crit = torch.nn.CrossEntropyLoss(weight=w, ignore_index=-1, label_smoothing=0.1)
        out = model(text, audio, mask)
        loss = crit(out.permute(0, 2, 1), emo)
        loss.backward()
        store_inputs = ragged != "rngfree"
        blob = {"text": text.numpy() if store_inputs else np.zeros(0), "audio": audio.numpy() if store_inputs else np.zeros(0),
                "mask": mask.numpy(), "emotion": emo.numpy(),
                "logits": out.detach().numpy(), "loss": np.array(loss.item()),
                "ce_weight": (w.numpy() if w is not None else np.zeros(0))}
        if name == "meld_base_2l":  # error level of an ideal bf16-operand pipeline on this case (the bf16 mode's yardstick)
            from oracle import bf16_emulation as E
            lg_e, _, g_e = E.loss_and_grads(params, cfg, text, audio, mask, emo, w)
            valid = ~mask
            ref_g = {O.canonical_name(n): p_.grad for n, p_ in model.named_parameters()}
            names = list(g_e)
            blob["bf16emu/logits"] = np.array(O.rel_l2(lg_e[valid], out.detach()[valid]))
            blob["bf16emu/grads"] = np.array(O.rel_l2(torch.cat([g_e[k].flatten() for k in names]), torch.cat([ref_g[k].flatten() for k in names])))
            print("   ideal bf16-operand pipeline: logits", float(blob["bf16emu/logits"]), "grads", float(blob["bf16emu/grads"]))
        seen = set()
        for pname, prm in model.named_parameters():  # named_parameters de-duplicates the shared norm
            cn = O.canonical_name(pname)
            assert cn not in seen
            seen.add(cn)
            for k, v in summarise(prm.grad).items():
                blob[f"grad/{cn}/{k}"] = v
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **blob)
        print(f"{name}: loss {loss.item():.12f}, {len(seen)} grads, logits {tuple(out.shape)}")


if __name__ == "__main__":
    main()
