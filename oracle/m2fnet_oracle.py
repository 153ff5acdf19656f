"""CPU oracle for the M2FNet hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference`
legs of `bench.py` may import this module.  The product path
(`multimodal-emotion-recognition_b200`) never imports it and has no CPU fallback.

What it is: a plain-tensor restatement (matmul / exp / sum only -- no `torch.nn`
modules, no `F.multi_head_attention_forward`, no `nn.TransformerEncoder`) of the
arithmetic the reference executes for

    M2FNet.forward                        /root/reference/src/model.py:102-145
    FusionAttentionModule.forward         /root/reference/src/model.py:13-20
    CrossEntropyLoss(weight?, ignore_index=-1, label_smoothing=0.1)
                                          /root/reference/src/train.py:48-50,229

The arithmetic itself lives in a third-party dependency of the reference, PyTorch
(pinned torch==1.13.1 in /root/reference/requirements.txt:11; 2.11.0 installed here):
`nn.TransformerEncoderLayer` post-norm branch (torch/nn/modules/transformer.py:951-982),
`F.multi_head_attention_forward` (torch/nn/functional.py:5833-5873, 6607-6690),
`nn.LayerNorm` (eps 1e-5, biased variance).  Formulas restated from SURVEY.md A.2.

Parity pinning: the reference ships no tests / golden vectors, so
the oracle is pinned against outputs of the reference module itself, imported in the
build container by `oracle/gen_golden.py` and committed under `tests/golden/`
(`tests/test_oracle_golden.py` checks every case, fp64, logits + loss + all gradients).

Gradients come from autograd over this restatement (fp64); the CUDA kernels' explicit
backward formulas are checked against it op by op in tests/test_gpu_ops.py.

Dropout: the reference draws 63 Bernoulli masks per forward from torch's global RNG
(SURVEY.md G9), which no other implementation can reproduce bit-for-bit.  The oracle
therefore takes the masks as *inputs* (`drop_masks[site_id] -> {0,1} tensor`) and applies
torch semantics `x * mask / (1 - p)`; site ids are defined in `site_id()` below and
mirrored by the CUDA library (`include/m2f_b200.h`, M2F_SITE_*).
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Optional

import torch

AUDIO, TEXT = 0, 1
MOD_NAME = {AUDIO: "audio", TEXT: "text"}

# dropout site ids ---------------------------------------------------------------------
# encoder layer (mod, t, l): k = 0 attention-probability dropout, 1 dropout1 (after the
# self-attention out-proj), 2 FFN dropout (after ReLU), 3 dropout2 (after linear2)
# modality level (mod): 0xF00 + {0: dropout before proj, 1: dropout after proj}
# fusion layer i: 0x3000 + 0x10*i + {0: attention-probability dropout, 1: dropout after the layer}
# classifier: 0x4000


def site_enc(mod: int, t: int, l: int, k: int) -> int:
    return 0x1000 * (mod + 1) + 0x100 * t + 0x10 * l + k


def site_mod(mod: int, k: int) -> int:
    return 0x1000 * (mod + 1) + 0xF00 + k


def site_fam(i: int, k: int) -> int:
    return 0x3000 + 0x10 * i + k


SITE_CLS = 0x4000


class Cfg(dict):
    """Attribute-access dict (what `Munch` gives the reference, src/utils.py:8-13)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    @classmethod
    def wrap(cls, d):
        if isinstance(d, dict):
            return cls({k: cls.wrap(v) for k, v in d.items()})
        return d


def default_model_cfg(**over) -> Cfg:
    """`model:` block of /root/reference/src/config.yaml:31-56."""
    c = {
        "dropout": 0.4,
        "AUDIO": {"enabled": True, "embedding_size": 768, "n_head": 8, "n_transformers": 1, "n_encoder_layers": 6},
        "TEXT": {"enabled": True, "embedding_size": 768, "n_head": 8, "n_transformers": 1, "n_encoder_layers": 6},
        "FAM": {"enabled": True, "embedding_size": 768, "n_head": 8, "n_layers": 5},
        "CLASSIFIER": {"hidden_size": 768, "output_size": 7, "n_layers": 2},
    }
    for k, v in over.items():
        if isinstance(v, dict):
            c[k].update(v)
        else:
            c[k] = v
    return Cfg.wrap(c)


FFN_DIM = 2048  # nn.TransformerEncoderLayer default dim_feedforward (model.py:61,73 pass none)
LN_EPS = 1e-5


def classifier_linear_indices(n_layers: int):
    """Indices of the nn.Linear slots inside `output_layer` (model.py:89-100)."""
    idx = [0]
    k = 1
    for _ in range(max(0, n_layers - 2)):
        idx.append(k + 1)  # ReLU at k, Linear at k+1
        k += 2
    idx.append(k + 2)  # ReLU at k, Dropout at k+1, Linear at k+2
    return idx


def param_shapes(cfg) -> Dict[str, tuple]:
    """state_dict names and shapes of the reference module.

    The shared final norm is listed once per `t`, as the reference
    state_dict does; `canonical_name` maps the aliases onto the `t = 0` entry.
    """
    out: Dict[str, tuple] = {}
    Df = cfg.FAM.embedding_size
    both = cfg.AUDIO.enabled and cfg.TEXT.enabled
    for mod, mc in ((AUDIO, cfg.AUDIO), (TEXT, cfg.TEXT)):
        if not mc.enabled:
            continue
        m, D = MOD_NAME[mod], mc.embedding_size
        for t in range(mc.n_transformers):
            for l in range(mc.n_encoder_layers):
                P = f"{m}_encoders.{t}.layers.{l}."
                out[P + "self_attn.in_proj_weight"] = (3 * D, D)
                out[P + "self_attn.in_proj_bias"] = (3 * D,)
                out[P + "self_attn.out_proj.weight"] = (D, D)
                out[P + "self_attn.out_proj.bias"] = (D,)
                out[P + "linear1.weight"] = (FFN_DIM, D)
                out[P + "linear1.bias"] = (FFN_DIM,)
                out[P + "linear2.weight"] = (D, FFN_DIM)
                out[P + "linear2.bias"] = (D,)
                for n in ("norm1", "norm2"):
                    out[P + n + ".weight"] = (D,)
                    out[P + n + ".bias"] = (D,)
            out[f"{m}_encoders.{t}.norm.weight"] = (D,)
            out[f"{m}_encoders.{t}.norm.bias"] = (D,)
        out[f"{m}_proj.weight"] = (Df, D)
        out[f"{m}_proj.bias"] = (Df,)
    if cfg.FAM.enabled:
        for i in range(cfg.FAM.n_layers):
            P = f"fusion_layers.{i}."
            out[P + "multihead_attention.in_proj_weight"] = (3 * Df, Df)
            out[P + "multihead_attention.in_proj_bias"] = (3 * Df,)
            out[P + "multihead_attention.out_proj.weight"] = (Df, Df)
            out[P + "multihead_attention.out_proj.bias"] = (Df,)
            out[P + "linear.weight"] = (Df, 2 * Df)
            out[P + "linear.bias"] = (Df,)
    Hc, C = cfg.CLASSIFIER.hidden_size, cfg.CLASSIFIER.output_size
    idx = classifier_linear_indices(cfg.CLASSIFIER.n_layers)
    in_sz = 2 * Df if both else Df
    for j, k in enumerate(idx):
        o = C if j == len(idx) - 1 else Hc
        out[f"output_layer.{k}.weight"] = (o, in_sz)
        out[f"output_layer.{k}.bias"] = (o,)
        in_sz = Hc
    return out


def canonical_name(name: str) -> str:
    """`{m}_encoders.{t}.norm.*` are aliases of one tensor."""
    parts = name.split(".")
    if len(parts) == 4 and parts[0].endswith("_encoders") and parts[2] == "norm":
        parts[1] = "0"
        return ".".join(parts)
    return name


def synth_param(name: str, shape: tuple, scale: Optional[float] = None) -> torch.Tensor:
    """Deterministic, RNG-free fp64 parameter values (a sin-hash of name and index), so
    the golden fixtures need not store weights.  LayerNorm weights are centred on 1."""
    h = 0
    for ch in canonical_name(name):
        h = (h * 131 + ord(ch)) % 1000003
    n = 1
    for s in shape:
        n *= s
    i = torch.arange(n, dtype=torch.float64)
    v = torch.sin(i * (0.7390851 + (h % 97) * 1e-3) + h * 0.61803) + 0.37 * torch.cos(i * 2.2360679 + h)
    if scale is None:
        fan_in = shape[-1] if len(shape) == 2 else 1
        scale = 0.9 / math.sqrt(fan_in) if len(shape) == 2 else 0.05
    v = v * scale
    if name.endswith(".weight") and len(shape) == 1:
        v = 1.0 + v
    return v.reshape(shape)


def synth_params(cfg, dtype=torch.float64) -> Dict[str, torch.Tensor]:
    return {k: synth_param(k, s).to(dtype) for k, s in param_shapes(cfg).items()}


def synth_inputs(B, S, Dt, Da, C=7):
    """RNG-free MELD-shaped batch (sin-hash values with the statistics of SURVEY.md section 8(d); ragged lengths from a
    fixed pattern, dialogue 0 full, last dialogue of length 1) so that large golden cases need not store their inputs."""
    def field(n, d, a, b, amp, off):
        i = torch.arange(n * d, dtype=torch.float64)
        return (amp * (torch.sin(i * a + 0.3) + 0.41 * torch.cos(i * b)) + off).reshape(n, d)
    text = field(B * S, Dt, 0.9127, 2.7183, 0.63 / 0.765, 0.025).reshape(B, S, Dt)
    audio = field(B * S, Da, 1.3371, 0.5772, 0.233 / 0.765, 0.0).reshape(B, S, Da)
    L = torch.tensor([S if b == 0 else (1 if b == B - 1 and B > 1 else 2 + (7 * b) % max(1, S - 2)) for b in range(B)])
    pos = torch.arange(S)[None, :]
    mask = pos >= L[:, None]
    emotion = ((torch.arange(B * S) * 5 + 3) % 11 % C).reshape(B, S)
    return {"text": text.masked_fill(mask[..., None], 0.0).float(), "audio": audio.masked_fill(mask[..., None], 0.0).float(),
            "padding_mask": mask, "emotion": emotion.masked_fill(mask, -1)}


# ----------------------------------------------------------------------------- forward ops


def layer_norm(x, w, b):
    mu = x.mean(dim=-1, keepdim=True)
    xc = x - mu
    var = (xc * xc).mean(dim=-1, keepdim=True)
    return xc / torch.sqrt(var + LN_EPS) * w + b


def _drop(x, site, drop_masks, p):
    if drop_masks is None or p == 0.0:
        return x
    m = drop_masks(site, x.shape) if callable(drop_masks) else drop_masks[site]
    return x * m.to(x.dtype) / (1.0 - p)


def mha(q_in, k_in, v_in, W, b, Wo, bo, H, keymask, site=None, drop_masks=None, p=0.0):
    """q_in/k_in/v_in: [B,S,E]; keymask: [B,S] bool (True = padded key)."""
    B, S, E = q_in.shape
    hd = E // H
    q = q_in @ W[0:E].T + b[0:E]
    k = k_in @ W[E:2 * E].T + b[E:2 * E]
    v = v_in @ W[2 * E:3 * E].T + b[2 * E:3 * E]
    q = q.reshape(B, S, H, hd).permute(0, 2, 1, 3) / math.sqrt(hd)
    k = k.reshape(B, S, H, hd).permute(0, 2, 1, 3)
    v = v.reshape(B, S, H, hd).permute(0, 2, 1, 3)
    s = q @ k.transpose(-1, -2)  # [B,H,S,S]
    s = s.masked_fill(keymask[:, None, None, :], float("-inf"))
    s = s - s.max(dim=-1, keepdim=True).values
    e = torch.exp(s)
    P = e / e.sum(dim=-1, keepdim=True)
    P = _drop(P, site, drop_masks, p)
    o = (P @ v).permute(0, 2, 1, 3).reshape(B, S, E)
    return o @ Wo.T + bo


def forward(params: Dict[str, torch.Tensor], cfg, text, audio, mask,
            drop_masks=None, p: float = 0.0, relu_taps: Optional[dict] = None, relu_force=None) -> torch.Tensor:
    """M2FNet.forward restated (model.py:102-145).  text [B,S,Dt], audio [B,S,Da],
    mask [B,S] bool (True = padding).  `p = 0` or `drop_masks is None` => eval().
    `relu_taps` (a dict) receives, per ReLU site, the smallest |pre-activation| over the valid rows relative to the
    site's rms: how close the nearest ReLU gate is to flipping (SURVEY.md G15 -- a gate within the forward's rounding
    error of zero makes the gradients of two correct implementations differ by O(1 / sqrt(rows x width))).
    `relu_force = (tau, "open" | "closed")` evaluates the network with every gate whose |pre-activation| is below tau x the
    site's rms forced to pass / to block its gradient (the forward value moves by < tau x rms): the gradients of the two
    settings bracket what any correct implementation with forward error < tau can return (tests use the spread as the G15
    allowance on top of the fixed tolerance)."""
    g = lambda n: params[canonical_name(n)]
    dm = drop_masks

    def relu(z, site_name):
        if relu_taps is not None:
            zv = z.detach()[~mask]
            nz = zv.abs()[zv != 0]  # exact zeros were produced by an earlier ReLU (the text half of a fusion concat): no gate to flip
            relu_taps[site_name] = float(nz.min() / zv.pow(2).mean().sqrt())
        if relu_force is not None:
            tau, how = relu_force
            zd = z.detach()
            rms = zd[~mask].pow(2).mean().sqrt()
            risk = (zd.abs() < tau * rms) & (zd != 0)
            return torch.where(risk, z if how == "open" else z * 0.0, torch.relu(z))
        return torch.relu(z)

    def modality(x, mod, mc):
        m = MOD_NAME[mod]
        for t in range(mc.n_transformers):
            y = x
            for l in range(mc.n_encoder_layers):
                P = f"{m}_encoders.{t}.layers.{l}."
                sa = mha(y, y, y, g(P + "self_attn.in_proj_weight"), g(P + "self_attn.in_proj_bias"),
                         g(P + "self_attn.out_proj.weight"), g(P + "self_attn.out_proj.bias"),
                         mc.n_head, mask, site_enc(mod, t, l, 0), dm, p)
                y = layer_norm(y + _drop(sa, site_enc(mod, t, l, 1), dm, p),
                               g(P + "norm1.weight"), g(P + "norm1.bias"))
                h = relu(y @ g(P + "linear1.weight").T + g(P + "linear1.bias"), P + "linear1")
                h = _drop(h, site_enc(mod, t, l, 2), dm, p)
                f = h @ g(P + "linear2.weight").T + g(P + "linear2.bias")
                y = layer_norm(y + _drop(f, site_enc(mod, t, l, 3), dm, p),
                               g(P + "norm2.weight"), g(P + "norm2.bias"))
            x = x + layer_norm(y, g(f"{m}_encoders.{t}.norm.weight"), g(f"{m}_encoders.{t}.norm.bias"))
        x = _drop(x, site_mod(mod, 0), dm, p)
        x = x @ g(f"{m}_proj.weight").T + g(f"{m}_proj.bias")
        return _drop(x, site_mod(mod, 1), dm, p)

    a = modality(audio, AUDIO, cfg.AUDIO) if cfg.AUDIO.enabled else None
    t_ = modality(text, TEXT, cfg.TEXT) if cfg.TEXT.enabled else None

    if cfg.FAM.enabled:
        for i in range(cfg.FAM.n_layers):
            P = f"fusion_layers.{i}."
            # query = text, key = audio, value = text (model.py:14)
            x = mha(t_, a, t_, g(P + "multihead_attention.in_proj_weight"),
                    g(P + "multihead_attention.in_proj_bias"),
                    g(P + "multihead_attention.out_proj.weight"),
                    g(P + "multihead_attention.out_proj.bias"),
                    cfg.FAM.n_head, mask, site_fam(i, 0), dm, p)
            u = relu(torch.cat((x, t_), dim=2), P + "cat")
            t_ = relu(u @ g(P + "linear.weight").T + g(P + "linear.bias"), P + "linear")
            t_ = _drop(t_, site_fam(i, 1), dm, p)
        z = torch.cat((a, t_), dim=2)
    elif cfg.AUDIO.enabled and cfg.TEXT.enabled:
        z = torch.cat((a, t_), dim=2)
    else:
        z = t_ if cfg.TEXT.enabled else a

    idx = classifier_linear_indices(cfg.CLASSIFIER.n_layers)
    for j, k in enumerate(idx):
        z = z @ g(f"output_layer.{k}.weight").T + g(f"output_layer.{k}.bias")
        if j < len(idx) - 1:
            z = relu(z, f"output_layer.{k}")
            if j == len(idx) - 2:
                z = _drop(z, SITE_CLS, dm, p)
    return z


def ce_loss(logits, target, weight=None, eps: float = 0.1, ignore_index: int = -1):
    """CrossEntropyLoss(weight, ignore_index=-1, label_smoothing=eps), mean reduction, on
    logits [B,S,C] / target [B,S] (the reference passes the (B,C,S) permuted view,
    train.py:229).  Closed form of SURVEY.md section 8(a) row a11."""
    C = logits.shape[-1]
    lg = logits.reshape(-1, C)
    tg = target.reshape(-1)
    valid = tg != ignore_index
    z = lg - lg.max(dim=-1, keepdim=True).values
    lp = z - torch.log(torch.exp(z).sum(dim=-1, keepdim=True))
    w = torch.ones(C, dtype=logits.dtype, device=logits.device) if weight is None else weight.to(logits.dtype)
    tsafe = torch.where(valid, tg, torch.zeros_like(tg))
    wy = w[tsafe] * valid.to(logits.dtype)
    nll = -(lp.gather(1, tsafe[:, None])[:, 0]) * wy
    smooth = -(lp * w[None, :]).sum(dim=-1) * valid.to(logits.dtype)
    return ((1.0 - eps) * nll.sum() + (eps / C) * smooth.sum()) / wy.sum()


def loss_and_grads(params, cfg, text, audio, mask, target, weight=None,
                   drop_masks=None, p: float = 0.0, relu_force=None):
    """Returns (logits, loss, {canonical param name: grad}).  Shared-norm aliases
    accumulate into the t = 0 entry, as autograd does for the reference."""
    leaves = {}
    for k, v in params.items():
        ck = canonical_name(k)
        if ck not in leaves:
            leaves[ck] = params[ck].detach().clone().requires_grad_(True)
    logits = forward(leaves, cfg, text, audio, mask, drop_masks, p, relu_force=relu_force)
    loss = ce_loss(logits, target, weight)
    names = list(leaves.keys())
    grads = torch.autograd.grad(loss, [leaves[n] for n in names], allow_unused=True)
    gd = {n: (g if g is not None else torch.zeros_like(leaves[n])) for n, g in zip(names, grads)}
    return logits.detach(), loss.detach(), gd


# ----------------------------------------------------------------------------- metrics


def rel_l2(x, ref) -> float:
    x = x.double().flatten()
    ref = ref.double().flatten()
    d = torch.linalg.vector_norm(ref).item()
    return torch.linalg.vector_norm(x - ref).item() / (d if d > 0 else 1.0)


def batch_metrics(logits, target):
    """Accuracy and weighted-F1 of one batch as src/test.py:65-71 computes them
    (sklearn semantics restated: F1 per class present in y_true or y_pred, weighted by
    the support in y_true)."""
    pred = logits.argmax(dim=-1)
    m = target != -1
    y, yp = target[m].flatten(), pred[m].flatten()
    n = y.numel()
    acc = (y == yp).double().mean().item()
    f1w = 0.0
    for c in torch.unique(torch.cat((y, yp))).tolist():
        tp = ((y == c) & (yp == c)).sum().item()
        fp = ((y != c) & (yp == c)).sum().item()
        fn = ((y == c) & (yp != c)).sum().item()
        f1 = 0.0 if (2 * tp + fp + fn) == 0 else 2 * tp / (2 * tp + fp + fn)
        f1w += f1 * (tp + fn) / n
    return acc, f1w


# ----------------------------------------------------------------------------- synthetic data


def synth_batch(B, S, Dt, Da, C=7, seed=1234, ragged=False, device="cpu", audio_kind="wav2vec2"):
    """MELD-shaped synthetic batch laid out as `collate_fn` does (src/dataset.py:71-89):
    statistics from SURVEY.md section 8(d)."""
    g = torch.Generator().manual_seed(seed)
    text = torch.randn(B, S, Dt, generator=g) * 0.63 + 0.025
    if audio_kind == "mel":
        audio = torch.randn(B, S, Da, generator=g)
        audio = audio / audio.norm(dim=-1, keepdim=True)
    else:
        audio = torch.randn(B, S, Da, generator=g) * 0.233
    probs = torch.tensor([.47, .17, .07, .11, .12, .03, .03])[:C]
    probs = probs / probs.sum()
    emotion = torch.multinomial(probs, B * S, replacement=True, generator=g).reshape(B, S)
    if ragged:
        L = torch.exp(torch.randn(B, generator=g) * 0.6 + 2.1).round().clamp(1, S).long()
        L[0] = S
        if B > 1:
            L[B - 1] = 1
    else:
        L = torch.full((B,), S, dtype=torch.long)
    pos = torch.arange(S)[None, :]
    mask = pos >= L[:, None]
    text = text.masked_fill(mask[..., None], 0.0)
    audio = audio.masked_fill(mask[..., None], 0.0)
    emotion = emotion.masked_fill(mask, -1)
    return {"text": text.to(device), "audio": audio.to(device),
            "padding_mask": mask.to(device), "emotion": emotion.to(device)}
