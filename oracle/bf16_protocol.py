"""SURVEY.md 8c protocol for the reduced-precision modes, on the CPU -- TEST INFRASTRUCTURE (oracle), not product code.

Protocol: the reference's default init under torch.manual_seed(0) (imported reference module when /root/reference is
there, else the oracle's own restatement of the same init), every parameter perturbed by 0.02 * randn (de-cloned layers,
SURVEY.md G6), full default config (6 + 6 encoder layers, 5 fusion layers), B = 8, S = 33 ragged, dropout 0, fp64 truth.

Against that truth this script evaluates *idealised operand-rounding pipelines*: every matrix product of the path has
its operands rounded to a chosen number of bf16 planes (1 plane = bf16, 2 planes = 16 mantissa bits, 0 = exact) while
accumulation, LayerNorm, softmax, residual stream and loss stay fp64.  The four knobs are the four operand classes a
training step has: activations X (forward and the weight-gradient B operand), weights W, gradients dY (both backward
products' A operand) and attention internals.  The table it prints is what decides which mode the product may call
"bf16": the cheapest one whose logits and concatenated gradients are within BASELINE.json's 2e-2.

    python oracle/bf16_protocol.py            # prints the table, writes profiles/r2_bf16_protocol_cpu.json
"""
from __future__ import annotations

import json
import math
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import m2fnet_oracle as O  # noqa: E402


def planes_round(x: torch.Tensor, n: int) -> torch.Tensor:
    """Sum of the first n bf16 planes of x (n = 0: x itself)."""
    if n == 0:
        return x
    out = torch.zeros_like(x)
    r = x
    for _ in range(n):
        h = r.to(torch.float32).to(torch.bfloat16).to(x.dtype)
        out = out + h
        r = r - h
    return out


class Mode:
    def __init__(self, x=1, w=1, g=1, attn=None, gx=None, gw=None, per=None, name="", noise=0.0, noise_seed=0):
        """x / w / g: planes of activations / weights / gradients as GEMM operands; attn: planes inside the attention core
        (default: x forward, g backward); gx / gw: planes of the activations / weights as operands of the two backward
        products (default x / w); per: {GEMM class: (x, w)} overrides for the forward products of that class
        (classes: qkv wo ffn1 ffn2 proj fam_qkv fam_wo fam_lin cls)."""
        self.x, self.w, self.g = x, w, g
        self.attn = x if attn is None else attn
        self.gx = x if gx is None else gx
        self.gw = w if gw is None else gw
        self.per = per or {}
        # additive noise of relative size `noise` (x the rms of the output) on every Linear's forward output: a model of a
        # correct implementation whose forward carries that much rounding error (fp32 accumulation over K terms); used to
        # measure how sensitive the GRADIENTS of this ReLU network are to forward error (gate flips)
        self.noise = float(noise)
        self.gen = torch.Generator().manual_seed(noise_seed) if noise else None
        self.name = name or f"x{x}w{w}g{g}"


def make_ops(mode: Mode):
    class RM(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, w, cls):  # x [..., K], w [N, K]
            px, pw = mode.per.get(cls, (mode.x, mode.w))
            xr, wr = planes_round(x, px), planes_round(w, pw)
            # what the backward products read: the leading planes of what the forward stored
            sx = min(px, mode.gx) if (px and mode.gx) else (px or mode.gx)
            sw = min(pw, mode.gw) if (pw and mode.gw) else (pw or mode.gw)
            ctx.save_for_backward(planes_round(x, sx), planes_round(w, sw))
            y = xr @ wr.T
            if mode.noise:
                y = y + mode.noise * y.pow(2).mean().sqrt() * torch.randn(y.shape, generator=mode.gen, dtype=y.dtype)
            return y

        @staticmethod
        def backward(ctx, g):
            xs, wr = ctx.saved_tensors
            gr = planes_round(g, mode.g)
            return gr @ wr, gr.reshape(-1, gr.shape[-1]).T @ xs.reshape(-1, xs.shape[-1]), None

    class RB(torch.autograd.Function):
        @staticmethod
        def forward(ctx, a, b):
            ar, br = planes_round(a, mode.attn), planes_round(b, mode.attn)
            ctx.save_for_backward(ar, br)
            return ar @ br

        @staticmethod
        def backward(ctx, g):
            ar, br = ctx.saved_tensors
            gr = planes_round(g, mode.g if mode.attn else 0)
            return gr @ br.transpose(-1, -2), ar.transpose(-1, -2) @ gr

    return (lambda x, w, cls: RM.apply(x, w, cls)), RB.apply


def forward(params, cfg, text, audio, mask, mode: Mode):
    mm, bmm = make_ops(mode)
    g = lambda n: params[O.canonical_name(n)]

    def mha(q_in, k_in, v_in, W, b, Wo, bo, H, kind=""):
        B, S, E = q_in.shape
        hd = E // H
        q = mm(q_in, W[0:E], kind + "qkv") + b[0:E]
        k = mm(k_in, W[E:2 * E], kind + "qkv") + b[E:2 * E]
        v = mm(v_in, W[2 * E:], kind + "qkv") + b[2 * E:]
        q = q.reshape(B, S, H, hd).permute(0, 2, 1, 3)
        k = k.reshape(B, S, H, hd).permute(0, 2, 1, 3)
        v = v.reshape(B, S, H, hd).permute(0, 2, 1, 3)
        s = bmm(q, k.transpose(-1, -2)) / math.sqrt(hd)
        s = s.masked_fill(mask[:, None, None, :], float("-inf"))
        P = torch.softmax(s, -1)
        o = bmm(P, v).permute(0, 2, 1, 3).reshape(B, S, E)
        return mm(o, Wo, kind + "wo") + bo

    def modality(x, mod, mc):
        m = O.MOD_NAME[mod]
        for t in range(mc.n_transformers):
            y = x
            for l in range(mc.n_encoder_layers):
                P = f"{m}_encoders.{t}.layers.{l}."
                sa = mha(y, y, y, g(P + "self_attn.in_proj_weight"), g(P + "self_attn.in_proj_bias"), g(P + "self_attn.out_proj.weight"),
                         g(P + "self_attn.out_proj.bias"), mc.n_head)
                y = O.layer_norm(y + sa, g(P + "norm1.weight"), g(P + "norm1.bias"))
                h = torch.relu(mm(y, g(P + "linear1.weight"), "ffn1") + g(P + "linear1.bias"))
                f = mm(h, g(P + "linear2.weight"), "ffn2") + g(P + "linear2.bias")
                y = O.layer_norm(y + f, g(P + "norm2.weight"), g(P + "norm2.bias"))
            x = x + O.layer_norm(y, g(f"{m}_encoders.{t}.norm.weight"), g(f"{m}_encoders.{t}.norm.bias"))
        return mm(x, g(f"{m}_proj.weight"), "proj") + g(f"{m}_proj.bias")

    a = modality(audio, 0, cfg.AUDIO)
    t_ = modality(text, 1, cfg.TEXT)
    for i in range(cfg.FAM.n_layers):
        P = f"fusion_layers.{i}."
        x = mha(t_, a, t_, g(P + "multihead_attention.in_proj_weight"), g(P + "multihead_attention.in_proj_bias"),
                g(P + "multihead_attention.out_proj.weight"), g(P + "multihead_attention.out_proj.bias"), cfg.FAM.n_head, "fam_")
        u = torch.relu(torch.cat((x, t_), 2))
        t_ = torch.relu(mm(u, g(P + "linear.weight"), "fam_lin") + g(P + "linear.bias"))
    z = torch.cat((a, t_), 2)
    idx = O.classifier_linear_indices(cfg.CLASSIFIER.n_layers)
    for j, k in enumerate(idx):
        z = mm(z, g(f"output_layer.{k}.weight"), "cls") + g(f"output_layer.{k}.bias")
        if j < len(idx) - 1:
            z = torch.relu(z)
    return z


def protocol_params(cfg, seed=0, perturb=0.02):
    """Default init of the reference under manual_seed(seed) + perturb * randn, as fp64 tensors keyed by state_dict name."""
    ref_path = "/root/reference/src/model.py"
    torch.manual_seed(seed)
    if os.path.exists(ref_path):
        import importlib.util
        import warnings
        warnings.filterwarnings("ignore")
        spec = importlib.util.spec_from_file_location("ref_model", ref_path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sd = {k: v.double() for k, v in mod.M2FNet(cfg).state_dict().items()}
    else:
        sd = default_init(cfg)
    gen = torch.Generator().manual_seed(seed + 1)
    out = {}
    for k, v in sd.items():
        ck = O.canonical_name(k)
        if ck not in out:
            out[ck] = v + perturb * torch.randn(v.shape, generator=gen, dtype=torch.float64)
    return {k: out[O.canonical_name(k)] for k in sd}


def default_init(cfg):
    """Restatement of the reference modules' default init distributions (SURVEY.md A.3 'Init'); used where the reference
    cannot be imported (GPU box)."""
    sd = {}
    template = {}
    for name, shape in O.param_shapes(cfg).items():
        parts = name.split(".")
        key = None
        if parts[0].endswith("_encoders") and parts[2] == "layers":
            key = parts[0] + "." + ".".join(parts[4:])
            if key in template:
                sd[name] = template[key].clone()
                continue
        p = torch.empty(shape, dtype=torch.float64)
        if name.endswith("in_proj_weight"):
            torch.nn.init.xavier_uniform_(p)
        elif name.endswith("in_proj_bias") or name.endswith("out_proj.bias"):
            p.zero_()
        elif ".norm" in name and len(shape) == 1:
            p.fill_(1.0) if name.endswith("weight") else p.zero_()
        elif len(shape) == 2:
            torch.nn.init.kaiming_uniform_(p, a=math.sqrt(5))
        else:
            bound = 1.0 / math.sqrt(O.param_shapes(cfg)[name[:-4] + "weight"][1])
            p.uniform_(-bound, bound)
        sd[name] = p
        if key is not None:
            template[key] = p.clone()
    return sd


def evaluate(params, cfg, batch, mode: Mode, truth):
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items() if O.canonical_name(k) == k}
    logits = forward(leaves, cfg, batch["text"].double(), batch["audio"].double(), batch["padding_mask"], mode)
    loss = O.ce_loss(logits, batch["emotion"])
    names = list(leaves)
    grads = dict(zip(names, torch.autograd.grad(loss, [leaves[n] for n in names])))
    t_logits, t_loss, t_grads = truth
    valid = ~batch["padding_mask"]
    per = {n: O.rel_l2(grads[n], t_grads[n]) for n in names if float(t_grads[n].abs().max()) > 0}
    cat = O.rel_l2(torch.cat([grads[n].flatten() for n in names]), torch.cat([t_grads[n].flatten() for n in names]))
    worst = sorted(per.items(), key=lambda kv: -kv[1])[:5]
    over = sorted(n for n, e in per.items() if e > 2e-2)
    return {"mode": mode.name, "logits": O.rel_l2(logits.detach()[valid], t_logits[valid]), "loss": abs(loss.item() - t_loss.item()) / abs(t_loss.item()),
            "grads_cat": cat, "worst": worst, "n_over_2e-2": len(over), "n_tensors": len(per),
            "argmax_same": float((logits.detach().argmax(-1)[valid] == t_logits.argmax(-1)[valid]).double().mean())}


def main():
    torch.set_num_threads(os.cpu_count() or 8)
    cfg = O.default_model_cfg(dropout=0.0)
    quick = "--quick" in sys.argv
    if quick:
        cfg = O.default_model_cfg(dropout=0.0, AUDIO=dict(n_encoder_layers=2), TEXT=dict(n_encoder_layers=2), FAM=dict(n_layers=2))
    out = []
    for init in ("protocol", "synth"):
        params = protocol_params(cfg) if init == "protocol" else O.synth_params(cfg)
        batch = O.synth_batch(8, 33, 768, 768, seed=1234, ragged=True)
        truth = O.loss_and_grads(params, cfg, batch["text"].double(), batch["audio"].double(), batch["padding_mask"], batch["emotion"])
        two = {c: (2, 2) for c in ("qkv", "wo", "ffn1", "ffn2", "proj", "fam_qkv", "fam_wo", "fam_lin", "cls")}
        sub = lambda *cs: {c: (2, 2) for c in cs}
        modes = [Mode(1, 1, 1, name="x1 w1 g1 (plain bf16 operands)"),
                 Mode(1, 1, 0, name="x1 w1 g exact"),
                 Mode(2, 1, 1, name="x2 w1 g1"),
                 Mode(1, 2, 1, name="x1 w2 g1"),
                 Mode(0, 0, 1, name="x,w exact, g1"),
                 Mode(2, 2, 1, name="fwd x2 w2 | bwd g1 x2 w2"),
                 Mode(2, 2, 1, gx=1, gw=1, name="fwd x2 w2 | bwd g1 x1 w1  [bf16x2f]"),
                 Mode(2, 2, 1, gx=1, gw=1, attn=1, name="bf16x2f, attention core 1 plane"),
                 Mode(1, 1, 1, per=sub("ffn1", "fam_wo", "fam_lin", "cls"), name="2 planes only in front of ReLUs"),
                 Mode(1, 1, 1, per=sub("ffn1", "ffn2", "fam_wo", "fam_lin", "cls"), name="2 planes ffn1 ffn2 fam_wo fam_lin cls"),
                 Mode(1, 1, 1, per=sub("qkv", "wo", "ffn1", "ffn2", "proj"), name="2 planes in the encoders only"),
                 Mode(2, 2, 2, name="x2 w2 g2 (bf16x2)"),
                 Mode(3, 3, 3, name="x3 w3 g3 (fp32-parity)")]
        if init == "synth":
            modes = [modes[0], modes[6], modes[-2]]
        for m in modes:
            r = evaluate(params, cfg, batch, m, truth)
            r["init"] = init
            out.append(r)
            print(f"[{init}] {r['mode']:34s} logits {r['logits']:.2e} loss {r['loss']:.1e} grads {r['grads_cat']:.2e} "
                  f"tensors>2e-2: {r['n_over_2e-2']}/{r['n_tensors']} worst {r['worst'][0][0]} {r['worst'][0][1]:.2e} argmax {r['argmax_same']:.3f}",
                  flush=True)
    if not quick:
        p = os.path.join(os.path.dirname(HERE), "profiles", "r2_bf16_protocol_cpu.json")
        with open(p, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
