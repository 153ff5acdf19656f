"""Parity of the product path on the other BASELINE.json configurations, checked against the CPU oracle run in fp64 on
the same inputs (the oracle itself is pinned to the reference by tests/test_oracle_golden.py; fixtures of these sizes
would be hundreds of MB, so the oracle is evaluated in the test):

* C3 dims: text 1024 / 16 heads (hd 64), audio 300 / 10 heads (hd 30 -> generic attention kernel), FAM 1024 / 16 heads,
  2 transformers sharing their final norm, 3-layer classifier head;
* C5 shape: S = 256 utterances per dialogue in eval / inference mode (attention beyond the S <= 64 fast path);
* the reference's training loop body (src/train.py:222-233) for several steps with the stock torch criterion and
  torch.optim.Adam: parameters after N steps must track the oracle trained the same way."""
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


def _build(P, cfg, planes):
    params = O.synth_params(cfg)
    model = P.M2FNet(cfg, planes=planes)
    model.load_state_dict({k: v.float() for k, v in params.items()}, strict=True)
    return model.to("cuda"), params


def test_c3_dims_fp32_parity(P):
    cfg = O.default_model_cfg(dropout=0.0,
                              TEXT=dict(embedding_size=1024, n_head=16, n_transformers=2, n_encoder_layers=1),
                              AUDIO=dict(embedding_size=300, n_head=10, n_transformers=2, n_encoder_layers=1),
                              FAM=dict(embedding_size=1024, n_head=16, n_layers=2),
                              CLASSIFIER=dict(hidden_size=1024, output_size=7, n_layers=3))
    B, S = 4, 9
    model, params = _build(P, cfg, planes=3)
    model.train()
    b = O.synth_inputs(B, S, 1024, 300)
    ref_logits, ref_loss, ref_grads = O.loss_and_grads(params, cfg, b["text"].double(), b["audio"].double(), b["padding_mask"], b["emotion"])
    out = model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda())
    loss = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)(out.permute(0, 2, 1), b["emotion"].cuda())
    loss.backward()
    valid = ~b["padding_mask"]
    assert O.rel_l2(out.detach().cpu()[valid], ref_logits[valid]) < 1e-4
    assert abs(loss.item() - ref_loss.item()) / abs(ref_loss.item()) < 1e-4
    got = {O.canonical_name(n): p.grad.detach().cpu() for n, p in model.named_parameters()}
    num = sum(float(((got[n].double() - g) ** 2).sum()) for n, g in ref_grads.items())
    den = sum(float((g ** 2).sum()) for g in ref_grads.values())
    assert (num / den) ** 0.5 < 1e-4
    worst = max((O.rel_l2(got[n], g), n) for n, g in ref_grads.items() if float(g.abs().max()) > 0)
    assert worst[0] < 1e-3, worst


@pytest.mark.parametrize("planes,tol", [(3, 1e-4), (1, 2e-2)])
def test_c5_long_dialogues_inference(P, planes, tol):
    cfg = O.default_model_cfg(dropout=0.4, TEXT=dict(n_encoder_layers=2), AUDIO=dict(n_encoder_layers=2), FAM=dict(n_layers=2))
    B, S = 3, 256
    model, params = _build(P, cfg, planes=planes)
    model.eval()
    b = O.synth_inputs(B, S, 768, 768)
    with torch.inference_mode():
        out = model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda())
        ref = O.forward(params, cfg, b["text"].double(), b["audio"].double(), b["padding_mask"])
    valid = ~b["padding_mask"]
    assert O.rel_l2(out.cpu()[valid], ref[valid]) < tol
    if planes == 3:
        assert bool((out.cpu().argmax(-1)[valid] == ref.argmax(-1)[valid]).all())


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16x2f", 2e-2), ("bf16", 2e-2)])
def test_inference_batch_of_many_tiles(P, precision, tol):
    """An inference batch large enough (52 dialogues x 256 utterances = 13 312 rows: 156 and more 256 x 256 tiles per GEMM) for the
    whole-model forward to run on the persistent two-accumulator GEMM (gemm_persist.cu) and, for one plane, the persistent
    tcgen05 attention with several (dialogue, head) items per CTA (416 items) -- against the fp64 oracle on the CPU, ragged lengths."""
    cfg = O.default_model_cfg(dropout=0.4, TEXT=dict(n_encoder_layers=1), AUDIO=dict(n_encoder_layers=1), FAM=dict(n_layers=1))
    B, S = 52, 256
    params = O.synth_params(cfg)
    model = P.M2FNet(cfg, precision=precision)
    model.load_state_dict({k: v.float() for k, v in params.items()}, strict=True)
    model = model.to("cuda").eval()
    b = O.synth_inputs(B, S, 768, 768)
    with torch.inference_mode():
        out = model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda())
        ref = O.forward(params, cfg, b["text"].double(), b["audio"].double(), b["padding_mask"])
    valid = ~b["padding_mask"]
    assert O.rel_l2(out.cpu()[valid], ref[valid]) < tol
    if precision == "fp32":
        assert bool((out.cpu().argmax(-1)[valid] == ref.argmax(-1)[valid]).all())


@pytest.mark.parametrize("B,S", [(1, 1), (3, 1), (2, 2), (2, 16), (3, 17), (2, 32), (2, 48), (3, 49), (2, 64), (2, 65), (2, 130), (2, 256), (1, 300)])
def test_sequence_length_edges(P, B, S):
    """Dialogue lengths at the edges of every attention tile size (single-utterance dialogues, the 16 / 32 / 48 / 64-row tiles of
    the short-dialogue kernels, the hand-over to the tiled long-dialogue kernels at 65, the tcgen05 inference kernel's limit of 256
    and beyond) with head sizes
    64 / 96 / 128: training forward + backward and the inference forward against the fp64 oracle at the fp32-parity tolerance."""
    cfg = O.default_model_cfg(dropout=0.0,
                              TEXT=dict(embedding_size=128, n_head=2, n_encoder_layers=1),
                              AUDIO=dict(embedding_size=96, n_head=1, n_encoder_layers=1),
                              FAM=dict(embedding_size=128, n_head=1, n_layers=1),
                              CLASSIFIER=dict(hidden_size=32, output_size=7, n_layers=2))
    model, params = _build(P, cfg, planes=3)
    model.train()
    b = O.synth_inputs(B, S, 128, 96)
    ref_logits, ref_loss, ref_grads = O.loss_and_grads(params, cfg, b["text"].double(), b["audio"].double(), b["padding_mask"], b["emotion"])
    out = model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda())
    loss = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)(out.permute(0, 2, 1), b["emotion"].cuda())
    loss.backward()
    valid = ~b["padding_mask"]
    assert O.rel_l2(out.detach().cpu()[valid], ref_logits[valid]) < 1e-4
    assert abs(loss.item() - ref_loss.item()) / abs(ref_loss.item()) < 1e-4
    got = {O.canonical_name(n): p.grad.detach().cpu() for n, p in model.named_parameters()}
    num = sum(float(((got[n].double() - g) ** 2).sum()) for n, g in ref_grads.items())
    den = sum(float((g ** 2).sum()) for g in ref_grads.values())
    assert (num / den) ** 0.5 < 1e-4
    model.eval()
    with torch.inference_mode():
        out_eval = model(b["text"].cuda(), b["audio"].cuda(), b["padding_mask"].cuda())
    assert O.rel_l2(out_eval.cpu()[valid], ref_logits[valid]) < 1e-4


def test_reference_train_loop_body_tracks_oracle(P):
    """Four iterations of the reference's step (zero_grad, forward, CE, backward, Adam; src/train.py:222-233) with the
    stock torch criterion and torch.optim.Adam on the product model, against the oracle trained by the same optimizer
    on the CPU in fp64 (dropout 0 so that both see the same function)."""
    cfg = O.default_model_cfg(dropout=0.0, TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=2),
                              AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=2), FAM=dict(embedding_size=64, n_head=4, n_layers=2),
                              CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2))
    B, S, steps = 6, 11, 4
    model, params = _build(P, cfg, planes=3)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=0.01)
    crit = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)
    leaves = {}
    for k in params:
        ck = O.canonical_name(k)
        if ck not in leaves:
            leaves[ck] = params[ck].detach().clone().requires_grad_(True)
    ref_opt = torch.optim.Adam(list(leaves.values()), lr=1e-3, weight_decay=0.01)
    b = O.synth_inputs(B, S, 128, 64)
    dev = {k: v.cuda() for k, v in b.items()}
    losses, ref_losses = [], []
    for _ in range(steps):
        opt.zero_grad()
        loss = crit(model(dev["text"], dev["audio"], dev["padding_mask"]).permute(0, 2, 1), dev["emotion"])
        loss.backward()
        opt.step()
        losses.append(loss.item())
        ref_opt.zero_grad()
        rl = O.ce_loss(O.forward(leaves, cfg, b["text"].double(), b["audio"].double(), b["padding_mask"]), b["emotion"])
        rl.backward()
        ref_opt.step()
        ref_losses.append(rl.item())
    assert losses[-1] < losses[0]
    for a, r in zip(losses, ref_losses):
        assert abs(a - r) / abs(r) < 1e-4, (losses, ref_losses)
    got = {O.canonical_name(n): p.detach().cpu().double() for n, p in model.named_parameters()}
    num = sum(float(((got[n] - v.detach()) ** 2).sum()) for n, v in leaves.items())
    den = sum(float((v.detach() ** 2).sum()) for v in leaves.values())
    assert (num / den) ** 0.5 < 1e-5


def test_fused_adam_keeps_the_weight_shadow_current(P):
    """FusedAdam updates the weights and re-splits their bf16 shadow in one pass (m2f_model_adam_step): the next forward
    must see exactly what an explicit shadow refresh of the updated parameters gives, and the update itself must equal
    torch.optim.Adam's."""
    cfg = O.default_model_cfg(dropout=0.0, TEXT=dict(embedding_size=128, n_head=4, n_encoder_layers=1),
                              AUDIO=dict(embedding_size=64, n_head=4, n_encoder_layers=1), FAM=dict(embedding_size=64, n_head=4, n_layers=2),
                              CLASSIFIER=dict(hidden_size=64, output_size=7, n_layers=2))
    b = {k: v.cuda() for k, v in O.synth_inputs(5, 7, 128, 64).items()}
    crit = torch.nn.CrossEntropyLoss(ignore_index=-1, label_smoothing=0.1)
    model, _ = _build(P, cfg, planes=1)
    twin, _ = _build(P, cfg, planes=1)
    model.train(), twin.train()
    opt = P.FusedAdam(model, lr=1e-3, weight_decay=0.01)
    ref = torch.optim.Adam(twin.parameters(), lr=1e-3, weight_decay=0.01)
    for _ in range(3):
        for m, o in ((model, opt), (twin, ref)):
            o.zero_grad()
            crit(m(b["text"], b["audio"], b["padding_mask"]).permute(0, 2, 1), b["emotion"]).backward()
            o.step()
    assert model._shadow_version is not None  # the fused path ran (shadow stayed current)
    rel = (torch.linalg.vector_norm(model.flat_parameters() - twin.flat_parameters()) / torch.linalg.vector_norm(twin.flat_parameters())).item()
    assert rel < 1e-6, rel
    model.eval()
    with torch.inference_mode():
        fused = model(b["text"], b["audio"], b["padding_mask"]).clone()
        model._shadow_version = None  # force m2f_model_refresh_shadow
        fresh = model(b["text"], b["audio"], b["padding_mask"])
    assert torch.equal(fused, fresh)
