"""Pins the oracle (oracle/m2fnet_oracle.py) against outputs of the unmodified reference
module recorded by oracle/gen_golden.py (tests/golden/*.npz): logits, loss and every
parameter gradient, fp64, tolerance 1e-10 relative."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import m2fnet_oracle as O
from oracle.gen_golden import CASES

TOL = 1e-10


def _cases(golden_dir=os.path.join(os.path.dirname(__file__), "golden")):
    names = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(golden_dir, "*.npz")))
    return [n for n in names if n != "collate_ref"]  # the batch-assembly fixture has its own tests (test_pipeline_cpu.py)


def test_all_cases_have_fixtures():
    assert set(_cases()) == set(CASES.keys())


@pytest.mark.parametrize("name", sorted(CASES.keys()))
def test_oracle_matches_reference_fixture(name, golden_dir):
    over, B, S, ragged, weighted = CASES[name]
    cfg = O.default_model_cfg(dropout=0.0, **over)
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    params = O.synth_params(cfg)
    from tests._golden import case_inputs
    text, audio = case_inputs(z, cfg, B, S, ragged)
    mask, emo = torch.from_numpy(z["mask"]), torch.from_numpy(z["emotion"])
    w = torch.from_numpy(z["ce_weight"]) if z["ce_weight"].size else None
    logits, loss, grads = O.loss_and_grads(params, cfg, text, audio, mask, emo, weight=w)
    assert O.rel_l2(logits, torch.from_numpy(z["logits"])) < TOL
    assert abs(loss.item() - float(z["loss"])) < TOL * abs(float(z["loss"]))
    names = sorted({k.split("/")[1] for k in z.files if k.startswith("grad/")})
    assert set(names) == set(grads.keys())
    for n in names:
        g = grads[n].double().flatten()
        if f"grad/{n}/full" in z.files:
            ref = torch.from_numpy(z[f"grad/{n}/full"])
            assert O.rel_l2(g, ref) < TOL, n
        else:
            step = int(z[f"grad/{n}/step"])
            assert O.rel_l2(g[::step], torch.from_numpy(z[f"grad/{n}/sample"])) < TOL, n
            assert O.rel_l2(g[:32], torch.from_numpy(z[f"grad/{n}/head"])) < TOL, n
            l2 = float(z[f"grad/{n}/l2"])
            assert abs(torch.linalg.vector_norm(g).item() - l2) <= TOL * max(l2, 1e-30), n
            assert abs(g.sum().item() - float(z[f"grad/{n}/sum"])) <= 1e-8 * max(l2, 1e-30), n


def test_padded_values_do_not_reach_valid_rows():
    """values stored at padded positions never influence valid rows."""
    cfg = O.default_model_cfg(dropout=0.0, **CASES["both_fam"][0])
    params = O.synth_params(cfg)
    b = O.synth_batch(3, 6, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size, ragged=True, seed=5)
    t, a, m = b["text"].double(), b["audio"].double(), b["padding_mask"]
    l0 = O.forward(params, cfg, t, a, m)
    t2 = t.masked_fill(m[..., None], 7.0)
    a2 = a.masked_fill(m[..., None], -3.0)
    l1 = O.forward(params, cfg, t2, a2, m)
    assert torch.equal(l0[~m], l1[~m])


def test_ce_closed_form_equals_torch():
    g = torch.Generator().manual_seed(3)
    logits = torch.randn(4, 6, 7, generator=g, dtype=torch.float64)
    tgt = torch.randint(0, 7, (4, 6), generator=g)
    tgt[1, 3:] = -1
    w = torch.rand(7, generator=g, dtype=torch.float64) + 0.5
    for weight in (None, w):
        ref = torch.nn.CrossEntropyLoss(weight=weight, ignore_index=-1, label_smoothing=0.1)(logits.permute(0, 2, 1), tgt)
        assert abs(O.ce_loss(logits, tgt, weight).item() - ref.item()) < 1e-13


def test_batch_metrics_match_sklearn():
    from sklearn.metrics import accuracy_score, f1_score
    g = torch.Generator().manual_seed(9)
    logits = torch.randn(5, 8, 7, generator=g)
    tgt = torch.randint(0, 7, (5, 8), generator=g)
    tgt[2, 4:] = -1
    acc, f1 = O.batch_metrics(logits, tgt)
    m = tgt != -1
    y, yp = tgt[m].numpy(), logits.argmax(-1)[m].numpy()
    assert abs(acc - accuracy_score(y, yp)) < 1e-12
    assert abs(f1 - f1_score(y, yp, average="weighted")) < 1e-12


def test_protocol_default_init_matches_the_reference_distributions():
    """oracle/bf16_protocol.default_init (used on the GPU box, where /root/reference does not exist) restates the init of the
    reference's torch modules: same names / shapes, same per-tensor spread, same cloned-layer structure."""
    import os
    import pytest
    import torch
    from oracle import bf16_protocol as BP
    from oracle import m2fnet_oracle as O
    if not os.path.exists("/root/reference/src/model.py"):
        pytest.skip("reference not present (GPU box)")
    cfg = O.default_model_cfg(dropout=0.0, AUDIO=dict(n_encoder_layers=2), TEXT=dict(n_encoder_layers=2), FAM=dict(n_layers=2))
    torch.manual_seed(0)
    mine = BP.default_init(cfg)
    ref = BP.protocol_params(cfg, seed=0, perturb=0.0)  # the imported reference's state_dict
    assert sorted(mine) == sorted(ref)
    for k in mine:
        a, b = mine[k].double(), ref[k].double()
        assert a.shape == b.shape, k
        if b.std() == 0:  # zero biases, LayerNorm ones / zeros
            assert torch.equal(a, b), k
        elif a.numel() >= 512:
            assert abs(a.std() / b.std() - 1) < 0.08 and abs(a.abs().max() / b.abs().max() - 1) < 0.08, k
        else:  # a handful of values: same order of magnitude is all that can be said
            assert 0.2 < float(a.abs().max() / b.abs().max()) < 5.0, k
    assert torch.equal(mine["text_encoders.0.layers.0.linear1.weight"], mine["text_encoders.0.layers.1.linear1.weight"])
    assert not torch.equal(mine["text_encoders.0.layers.0.linear1.weight"], mine["audio_encoders.0.layers.0.linear1.weight"])
