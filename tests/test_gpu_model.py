"""Model-level parity on the GPU: the product `M2FNet` (one C call forward, one C call backward, sm_100a kernels)
against the golden fixtures recorded from the UNMODIFIED reference module in fp64 (tests/golden/, oracle/gen_golden.py).
Tolerances are BASELINE.json's: 1e-4 relative in the fp32-parity mode (planes = 3), 2e-2 in the bf16 training mode
(planes = 2 forward / 1 backward, "bf16x2f");
logits are compared on valid rows only; per-tensor gradient bounds follow SURVEY.md 8c."""
import pytest
import torch

from oracle import m2fnet_oracle as O
from oracle.gen_golden import CASES
from tests._golden import case_inputs, grad_errors, load_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    return pkg


def run_case(P, name, planes, fused_loss=False, planes_bwd=None):
    over, B, S, ragged, weighted = CASES[name]
    cfg = O.default_model_cfg(dropout=0.0, **over)
    z = load_case(name)
    model = P.M2FNet(cfg, planes=planes, planes_bwd=planes_bwd)
    model.load_state_dict({k: v.float() for k, v in O.synth_params(cfg).items()}, strict=True)
    model = model.to("cuda")
    model.train()
    text, audio = (t.float().cuda() for t in case_inputs(z, cfg, B, S, ragged))
    mask, emo = torch.from_numpy(z["mask"]).cuda(), torch.from_numpy(z["emotion"]).cuda()
    w = torch.from_numpy(z["ce_weight"]).float().cuda() if z["ce_weight"].size else None
    crit = (P.FusedCrossEntropyLoss(weight=w) if fused_loss
            else torch.nn.CrossEntropyLoss(weight=w, ignore_index=-1, label_smoothing=0.1))
    out = model(text, audio, mask)
    loss = crit(out.permute(0, 2, 1), emo)
    loss.backward()
    torch.cuda.synchronize()
    ref_logits = torch.from_numpy(z["logits"])
    valid = ~torch.from_numpy(z["mask"])
    e_logits = O.rel_l2(out.detach().cpu()[valid], ref_logits[valid])
    e_loss = abs(loss.item() - float(z["loss"])) / abs(float(z["loss"]))
    grads = {O.canonical_name(n): p.grad for n, p in model.named_parameters()}
    per, e_grad = grad_errors(z, grads)
    pred_same = (out.detach().cpu().argmax(-1)[valid] == ref_logits.argmax(-1)[valid]).float().mean().item()
    return e_logits, e_loss, e_grad, per, pred_same


@pytest.mark.parametrize("name", sorted(CASES))
def test_fp32_parity_mode(P, name):
    e_logits, e_loss, e_grad, per, pred_same = run_case(P, name, planes=3)
    worst = max(per.items(), key=lambda kv: kv[1])
    print(f"{name}: logits {e_logits:.2e} loss {e_loss:.2e} grads {e_grad:.2e} worst {worst}")
    assert e_logits < 1e-4 and e_loss < 1e-4 and e_grad < 1e-4
    assert worst[1] < 1e-3, worst
    assert pred_same == 1.0


@pytest.mark.parametrize("name", sorted(CASES))
def test_bf16_training_mode_meets_2e2(P, name):
    """The declared bf16 training mode (precision "bf16x2f": two bf16 planes in the forward, one in the backward -- the
    default, and what bench.py's headline runs) at BASELINE.json's numbers: 2e-2 on logits, loss AND the concatenated
    gradients, every case.  SURVEY.md 8c's protocol (real init, full depth, per-tensor list) is tests/test_gpu_protocol.py."""
    over, B, S, ragged, weighted = CASES[name]
    e_logits, e_loss, e_grad, per, pred_same = run_case(P, name, planes=2, planes_bwd=1)
    worst = max(per.items(), key=lambda kv: kv[1])
    print(f"{name}: logits {e_logits:.2e} loss {e_loss:.2e} grads {e_grad:.2e} worst {worst}")
    assert e_logits < 2e-2 and e_loss < 2e-2 and e_grad < 2e-2
    assert pred_same == 1.0


@pytest.mark.parametrize("name", sorted(CASES))
def test_plain_bf16_mode(P, name):
    """One bf16 plane everywhere (precision "bf16"): the inference mode.  Logits and loss meet 2e-2; its gradients do NOT
    (ReLU gates decided by a rounded forward, profiles/r2_bf16_protocol_cpu.json) -- which is why it is not the training
    default -- and are only held to the level of an ideal bf16-operand pipeline (oracle/bf16_emulation.py)."""
    e_logits, e_loss, e_grad, per, pred_same = run_case(P, name, planes=1)
    worst = max(per.items(), key=lambda kv: kv[1])
    print(f"{name}: logits {e_logits:.2e} loss {e_loss:.2e} grads {e_grad:.2e} worst {worst}")
    assert e_logits < 2e-2 and e_loss < 2e-2
    if name == "meld_base_2l":
        z = load_case(name)
        assert e_logits < 1.25 * float(z["bf16emu/logits"]) and e_grad < 1.25 * float(z["bf16emu/grads"])
    else:
        assert e_grad < 0.3


def test_bf16x2_mode_meets_2e2_on_gradients(P):
    """Two bf16 planes everywhere (16 mantissa bits, 3 + 3 + 3 plane products per Linear)."""
    e_logits, e_loss, e_grad, per, pred_same = run_case(P, "meld_base_2l", planes=2)
    print(f"bf16x2: logits {e_logits:.2e} loss {e_loss:.2e} grads {e_grad:.2e}")
    assert e_logits < 1e-3 and e_grad < 2e-2 and pred_same == 1.0


def test_fused_loss_matches_torch_criterion(P):
    a = run_case(P, "shared_norm_cls3_weighted", planes=3, fused_loss=False)
    b = run_case(P, "shared_norm_cls3_weighted", planes=3, fused_loss=True)
    assert abs(a[1] - b[1]) < 1e-5 and b[2] < 1e-4


def test_eval_forward_equals_train_forward_at_p0(P):
    """with dropout 0, train() and eval() agree; the inference plan reuses buffers, results must not change."""
    over, B, S, ragged, weighted = CASES["both_fam"]
    cfg = O.default_model_cfg(dropout=0.0, **over)
    z = load_case("both_fam")
    model = P.M2FNet(cfg, planes=3)
    model.load_state_dict({k: v.float() for k, v in O.synth_params(cfg).items()}, strict=True)
    model = model.to("cuda")
    text, audio = torch.from_numpy(z["text"]).float().cuda(), torch.from_numpy(z["audio"]).float().cuda()
    mask = torch.from_numpy(z["mask"]).cuda()
    model.train()
    a = model(text, audio, mask).detach()
    model.eval()
    with torch.inference_mode():
        b = model(text, audio, mask)
    assert torch.equal(a, b)


def test_cpu_tensors_raise(P):
    cfg = O.default_model_cfg(dropout=0.0, **CASES["audio_only"][0])
    model = P.M2FNet(cfg)
    with pytest.raises(RuntimeError, match="no CPU path"):
        model(torch.zeros(1, 2, 32), torch.zeros(1, 2, 32), torch.zeros(1, 2, dtype=torch.bool))
