"""FusedAdam(in_backward=True): the optimizer update of every gradient range runs inside loss.backward() (m2f_plan_set_optimizer),
overlapped with the rest of the backward.  It must be the SAME update: parameters, Adam moments and the bf16 weight shadow after
several steps are bit-identical to the ordinary path (backward, then FusedAdam.step()), with dropout on, in the graph and the
eager execution modes.  Reference semantics: torch.optim.Adam(lr, weight_decay) as src/train.py:56,231 uses it."""
import pytest
import torch

from oracle import m2fnet_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg._lib.check(pkg._lib.lib().m2f_check_device(), "check_device")
    return pkg


def train(P, cfg, in_backward, steps, B, S, precision="bf16x2f", lr_schedule=None):
    params = O.synth_params(cfg)
    torch.manual_seed(5)
    model = P.M2FNet(cfg, precision=precision)
    model.load_state_dict({k: v.float() for k, v in params.items()}, strict=True)
    model = model.to("cuda").train()
    model._seed = 1234  # same dropout stream in both runs
    crit = P.FusedCrossEntropyLoss().cuda()
    opt = P.FusedAdam(model, lr=5e-3, weight_decay=0.01, in_backward=in_backward)
    losses = []
    for i in range(steps):
        b = O.synth_inputs(B, S, cfg.TEXT.embedding_size, cfg.AUDIO.embedding_size)
        if lr_schedule:
            opt.param_groups[0]["lr"] = lr_schedule[i]
        opt.zero_grad(set_to_none=True)
        out = model(b["text"].cuda() * (1 + 0.1 * i), b["audio"].cuda(), b["padding_mask"].cuda())
        loss = crit(out.permute(0, 2, 1), b["emotion"].cuda())
        loss.backward()
        opt.step()
        losses.append(loss.item())
    torch.cuda.synchronize()
    return model.flat_parameters().clone(), opt._m.clone(), opt._v.clone(), model._shadow.clone(), losses, model, opt


CFGS = {
    "both_fam": dict(AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=3), TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=3),
                     FAM=dict(embedding_size=64, n_head=4, n_layers=3), CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2)),
    "text_only": dict(AUDIO=dict(enabled=False), TEXT=dict(embedding_size=64, n_head=4, n_encoder_layers=2),
                      FAM=dict(enabled=False, embedding_size=64, n_head=4), CLASSIFIER=dict(hidden_size=32, output_size=7, n_layers=3)),
    "meld_dims": dict(AUDIO=dict(n_encoder_layers=2), TEXT=dict(n_encoder_layers=2), FAM=dict(n_layers=2)),
}


@pytest.mark.parametrize("name", sorted(CFGS))
@pytest.mark.parametrize("graph", ["1", "0"])
def test_update_inside_backward_is_bit_identical(P, name, graph, monkeypatch):
    monkeypatch.setenv("M2F_GRAPH", graph)
    cfg = O.default_model_cfg(dropout=0.4, **CFGS[name])
    B, S = (4, 33) if name == "meld_dims" else (5, 9)
    sched = [5e-3, 4e-3, 3e-3, 2e-3]
    a = train(P, cfg, False, 4, B, S, lr_schedule=sched)
    b = train(P, cfg, True, 4, B, S, lr_schedule=sched)
    assert a[4] == b[4], (a[4], b[4])  # losses of every step: the forward of step k saw the same weights
    for x, y, what in zip(a[:4], b[:4], ("parameters", "exp_avg", "exp_avg_sq", "bf16 shadow")):
        assert torch.equal(x, y), what
    assert a[6]._t == b[6]._t == 4


def test_misuse_is_an_error(P):
    cfg = O.default_model_cfg(dropout=0.0, **CFGS["both_fam"])
    *_, model, opt = train(P, cfg, True, 1, 3, 5)
    with pytest.raises(RuntimeError, match="no backward since the last step"):
        opt.step()
    b = O.synth_inputs(3, 5, 128, 64)
    crit = P.FusedCrossEntropyLoss().cuda()
    opt.zero_grad(set_to_none=True)
    crit(model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda()).permute(0, 2, 1), b["emotion"].cuda()).backward()
    with pytest.raises(RuntimeError, match="twice without step"):
        crit(model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda()).permute(0, 2, 1), b["emotion"].cuda()).backward()
    opt.step()
    opt.detach_from_backward()  # ordinary stepping again
    opt.zero_grad(set_to_none=True)
    crit(model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda()).permute(0, 2, 1), b["emotion"].cuda()).backward()
    before = model.flat_parameters().clone()
    torch.cuda.synchronize()
    assert torch.equal(before, model.flat_parameters())  # the backward left the parameters alone
    opt.step()
    torch.cuda.synchronize()
    assert not torch.equal(before, model.flat_parameters())
