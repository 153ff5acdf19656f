"""Host-side mirror of the reference's `model` module: `M2FNet` and `FusionAttentionModule`.

Same constructor (`M2FNet(config.model)`), same `forward(text, audio, mask)` signature, same `state_dict()` names
and shapes (incl. the shared final LayerNorm aliases) as /root/reference/src/model.py:5-145, so the
reference's `train.py` / `test.py` run unchanged with this module on `sys.path` as `model`.

What differs is what executes: every parameter is a view into ONE fp32 arena laid out by libm2f_b200
(`m2f_model_*`), the forward is ONE C call (`m2f_plan_forward`) and the backward is ONE C call
(`m2f_plan_backward`) into hand-written sm_100a kernels.  There is no eager / CPU fallback: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
import math
import operator
from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from . import _lib
from ._lib import ModalityCfg, ModelCfg

FFN_DIM = 2048  # nn.TransformerEncoderLayer's default dim_feedforward (src/model.py:61,73 pass none)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class _Node(nn.Module):
    """Name-space node: gives parameters the dotted names of the reference state_dict."""

    def forward(self, *a, **k):  # pragma: no cover - never called
        raise RuntimeError("structural node of M2FNet; call the model itself")


class FusionAttentionModule(_Node):
    """Name-space twin of src/model.py:5-20.  The arithmetic (MHA with query = text, key = audio, value = text,
    concat, ReLU, Linear, ReLU) runs inside m2f_plan_forward; this class only carries the parameters."""


def _cfg_get(c, key, default=None):
    try:
        return getattr(c, key)
    except AttributeError:
        if isinstance(c, dict) and key in c:
            return c[key]
        if default is not None:
            return default
        raise


# Precision modes: (planes of the forward's operands, planes of the backward's operands).
#   bf16x2f  (default) 16-bit-mantissa forward, plain bf16 backward: the cheapest mode that meets BASELINE.json's bf16
#            tolerance (2e-2) on logits AND gradients under SURVEY.md 8c's protocol (tests/test_gpu_protocol.py): what bf16
#            training loses is almost entirely ReLU gates / softmax weights decided differently by a rounded forward
#            (profiles/r2_bf16_protocol_cpu.json), so the forward gets the second plane and the backward does not.
#   bf16     plain bf16 operands everywhere (logits within 2e-2, gradients ~2.5e-2 under the same protocol): inference.
#   bf16x2   two planes everywhere;  fp32: three planes, the 1e-4 parity mode.
PRECISION_MODES = {"bf16x2f": (2, 1), "bf16": (1, 1), "1": (1, 1), "bf16x2": (2, 2), "2": (2, 2), "fp32": (3, 3), "bf16x3": (3, 3),
                   "3": (3, 3)}


def _default_planes():
    import os
    mode = os.environ.get("M2F_PRECISION", "bf16x2f").lower()
    if mode not in PRECISION_MODES:
        raise ValueError(f"M2F_PRECISION={mode!r}: expected one of {sorted(PRECISION_MODES)}")
    return PRECISION_MODES[mode]


_VERSION_OF = operator.attrgetter("_version")


class _Plan:
    """One (B, S, training) geometry: the C plan handle plus its torch-owned workspace."""

    def __init__(self, model: "M2FNet", B: int, S: int, training: bool):
        L = _lib.lib()
        h = C.c_void_p()
        _lib.check(L.m2f_plan_create(model._handle, B, S, int(training), C.byref(h)), "plan_create")
        self.handle = h
        self.B, self.S, self.training = B, S, training
        self.ws_bytes = int(L.m2f_plan_workspace_bytes(h))
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=model._arena.device)
        self.token = 0  # bumped by every forward; a backward must match the forward that filled the workspace
        if training and model._exchange is not None:
            model._attach_exchange(self)
        if training and model._tail_opt is not None:
            model._attach_optimizer(self)

    def destroy(self):
        if self.handle:
            _lib.lib().m2f_plan_destroy(self.handle)
            self.handle = None


class _M2FStep(torch.autograd.Function):
    """forward = m2f_plan_forward, backward = m2f_plan_backward.  Parameter gradients are written by the C call
    straight into the model's gradient arena and attached as `p.grad` views (no per-parameter autograd traffic);
    `anchor` is only there so that autograd records a node."""

    @staticmethod
    def forward(ctx, model, plan, text, audio, mask_u8, dropout_on, anchor):
        logits = model._launch_forward(plan, text, audio, mask_u8, dropout_on)
        ctx.model, ctx.plan, ctx.token = model, plan, plan.token
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        ctx.model._launch_backward(ctx.plan, ctx.token, dlogits)
        return None, None, None, None, None, None, None


class M2FNet(nn.Module):
    """Drop-in for the reference `M2FNet` (src/model.py:22-145)."""

    def __init__(self, config, planes: Optional[int] = None, planes_bwd: Optional[int] = None, precision: Optional[str] = None):
        """`precision`: one of PRECISION_MODES (default $M2F_PRECISION or "bf16x2f"); `planes` / `planes_bwd` set the two
        plane counts directly (planes alone = the same count forward and backward, as in round 1)."""
        super().__init__()
        self.config = config
        # the reference's constructor checks, same messages (src/model.py:32-35)
        a, t, f, c = config.AUDIO, config.TEXT, config.FAM, config.CLASSIFIER
        if not a.enabled and not t.enabled:
            raise ValueError("At least one of audio and text must be enabled!")
        if f.enabled and not (a.enabled and t.enabled):
            raise ValueError("Fusion Attention Module can only be used with both audio and text enabled!")
        for mc in (a, t):
            if mc.enabled:
                assert mc.embedding_size % mc.n_head == 0, "embed_dim must be divisible by num_heads"
        if f.enabled:
            assert f.embedding_size % f.n_head == 0, "embed_dim must be divisible by num_heads"
        if precision is not None:
            if precision.lower() not in PRECISION_MODES:
                raise ValueError(f"precision={precision!r}: expected one of {sorted(PRECISION_MODES)}")
            planes, planes_bwd = PRECISION_MODES[precision.lower()]
        if planes is None:
            planes = _cfg_get(config, "planes", 0) or None
        if planes is None:
            planes, dflt_bwd = _default_planes()
            planes_bwd = dflt_bwd if planes_bwd is None else planes_bwd
        self.planes = int(planes)
        self.planes_bwd = int(planes_bwd) if planes_bwd else self.planes
        if not 1 <= self.planes_bwd <= self.planes <= 3:
            raise ValueError(f"planes={self.planes}, planes_bwd={self.planes_bwd}: need 1 <= planes_bwd <= planes <= 3")
        cc = ModelCfg()
        for dst, src in ((cc.audio, a), (cc.text, t)):
            dst.enabled, dst.dim, dst.n_head = int(bool(src.enabled)), int(src.embedding_size), int(src.n_head)
            dst.n_transformers, dst.n_layers = int(src.n_transformers), int(src.n_encoder_layers)
        cc.fam_enabled, cc.fam_dim, cc.fam_heads, cc.fam_layers = int(bool(f.enabled)), int(f.embedding_size), int(f.n_head), int(f.n_layers)
        cc.cls_hidden, cc.cls_out, cc.cls_layers = int(c.hidden_size), int(c.output_size), int(c.n_layers)
        cc.dropout, cc.ffn_dim, cc.planes, cc.planes_bwd = float(config.dropout), FFN_DIM, self.planes, self.planes_bwd
        self._ccfg = cc
        L = _lib.lib()
        h = C.c_void_p()
        _lib.check(L.m2f_model_create(C.byref(cc), C.byref(h)), "model_create")
        self._handle = h
        self._arena_floats = int(L.m2f_model_arena_floats(h))
        self._off_1d = int(L.m2f_model_arena_1d_offset(h))
        self._shadow_bytes = int(L.m2f_model_shadow_bytes(h))
        self._layout: List[Tuple[str, int, Tuple[int, ...]]] = []
        n = int(L.m2f_model_num_params(h))
        buf = C.create_string_buffer(256)
        off, rows, cols = C.c_int64(), C.c_int64(), C.c_int64()
        for i in range(n):
            _lib.check(L.m2f_model_param_info(h, i, buf, 256, C.byref(off), C.byref(rows), C.byref(cols)), "param_info")
            shape = (rows.value, cols.value) if cols.value else (rows.value,)
            self._layout.append((buf.value.decode(), off.value, shape))
        # parameters = views into one arena, registered under the reference's dotted names
        self._arena = torch.zeros(self._arena_floats, dtype=torch.float32)
        self._flat: List[Tuple[nn.Parameter, int, int]] = []
        for name, o, shape in self._layout:
            p = nn.Parameter(self._arena[o:o + math.prod(shape)].view(shape))
            self._register_dotted(name, p)
            self._flat.append((p, o, math.prod(shape)))
        for m_name, mc in (("audio", a), ("text", t)):  # shared final norm: one module object under every t (G5)
            if mc.enabled:
                encs = getattr(self, f"{m_name}_encoders")
                for ti in range(1, mc.n_transformers):
                    getattr(encs, str(ti)).add_module("norm", getattr(encs, "0").norm)
        self.reset_parameters()
        self._plist = [p for p, _, _ in self._flat]
        self._anchor_box = [None]  # in a list: assigning a Parameter as an attribute would register it under that name
        self._grad_arena: Optional[torch.Tensor] = None
        self._exchange = None            # (PeerGroup, scale, ctas, bucket_floats): gradient exchange inside the backward
        self._tail_opt = None            # FusedAdam(overlap_exchange=True): the update of each exchanged range runs behind the backward
        self._grad_views = None          # p.grad views into the gradient arena, built once
        self._grads_cleared = False      # FusedAdam.zero_grad(): views still attached, contents dead
        self._shadow: Optional[torch.Tensor] = None
        self._shadow_version = None
        self._plans: Dict[Tuple[int, int, bool], _Plan] = {}
        self._seed = int(torch.initial_seed()) & ((1 << 64) - 1)
        self._step = 0
        self.launches = 0  # C-ABI model-level calls issued (bench.py reports kernel launches separately)
        # data-parallel hooks (parallel.GradBucketReducer): on_bucket(lo, hi) per finished gradient range, on_done()
        self._bucket_hook = None
        self._bucket_floats = 0
        self._done_hook = None

    # ------------------------------------------------------------------------------------------ structure
    def _register_dotted(self, name: str, p: nn.Parameter):
        parts = name.split(".")
        node: nn.Module = self
        for i, part in enumerate(parts[:-1]):
            child = node._modules.get(part)
            if child is None:
                is_fam = len(parts) > 2 and parts[0] == "fusion_layers" and i == 1
                child = FusionAttentionModule() if is_fam else _Node()
                node.add_module(part, child)
            node = child
        node.register_parameter(parts[-1], p)

    @torch.no_grad()
    def reset_parameters(self):
        """Same initial distributions as the reference's torch modules (SURVEY.md A.3 'Init'); the encoder layers of
        one modality start as clones of one template layer (G6)."""
        by_name = {n: p for n, p in self.named_parameters()}
        template: Dict[str, torch.Tensor] = {}
        for name, p in by_name.items():
            parts = name.split(".")
            key = None
            if parts[0].endswith("_encoders") and parts[2] == "layers":
                key = parts[0] + "." + ".".join(parts[4:])
                if key in template:
                    p.copy_(template[key])
                    continue
            if name.endswith("in_proj_weight"):
                nn.init.xavier_uniform_(p)
            elif name.endswith("in_proj_bias") or name.endswith("out_proj.bias"):
                p.zero_()
            elif ".norm" in name and p.dim() == 1:
                p.fill_(1.0) if name.endswith("weight") else p.zero_()
            elif p.dim() == 2:
                nn.init.kaiming_uniform_(p, a=math.sqrt(5))
            else:  # Linear bias: U(-1/sqrt(fan_in), 1/sqrt(fan_in))
                w = by_name[name[:-4] + "weight"]
                bound = 1.0 / math.sqrt(w.shape[1])
                p.uniform_(-bound, bound)
            if key is not None:
                template[key] = p.detach().clone()

    # ------------------------------------------------------------------------------------------ arena upkeep
    def _apply(self, fn, recurse=True):
        super()._apply(fn, recurse)
        self._repack()
        return self

    @torch.no_grad()
    def _repack(self):
        """Make every parameter a view of one contiguous fp32 arena again (after .to(device) re-allocated them)."""
        p0 = self._flat[0][0]
        dev = p0.device
        for p, _, _ in self._flat:
            if p.dtype != torch.float32:
                raise RuntimeError("M2FNet parameters must stay fp32 (master weights); precision is chosen with `planes`")
        arena = torch.empty(self._arena_floats, dtype=torch.float32, device=dev)
        arena.zero_()
        for p, o, n in self._flat:
            view = arena[o:o + n].view(p.shape)
            view.copy_(p.data)
            p.data = view
            p.grad = None
        self._arena = arena
        self._plist = [p for p, _, _ in self._flat]
        self._grad_arena = None
        self._grad_views = None
        self._grads_cleared = False
        self._shadow = None
        self._shadow_version = None
        for pl in self._plans.values():
            pl.destroy()
        self._plans.clear()

    def _views_intact(self) -> bool:
        base = self._arena.data_ptr()
        for idx in (0, len(self._flat) // 2, len(self._flat) - 1):
            p, o, _ = self._flat[idx]
            if p.data_ptr() != base + 4 * o:
                return False
        return True

    def _param_version(self) -> int:
        return sum(map(_VERSION_OF, self._plist))

    def _ensure_shadow(self):
        if not self._views_intact():
            self._repack()
        ver = self._param_version()
        if self._shadow is None:
            self._shadow = torch.empty(self._shadow_bytes, dtype=torch.uint8, device=self._arena.device)
            self._shadow_version = None
        if self._shadow_version != ver:
            _lib.check(_lib.lib().m2f_model_refresh_shadow(self._handle, self._arena.data_ptr(), self._shadow.data_ptr(), _stream()),
                       "refresh_shadow")
            self._shadow_version = ver

    def _plan(self, B: int, S: int, training: bool) -> _Plan:
        key = (B, S, training)
        pl = self._plans.get(key)
        if pl is None:
            pl = _Plan(self, B, S, training)
            self._plans[key] = pl
        return pl

    # ------------------------------------------------------------------------------------------ execution
    def _launch_forward(self, plan: _Plan, text, audio, mask_u8, dropout_on: bool) -> torch.Tensor:
        C_out = self._ccfg.cls_out
        logits = torch.empty((plan.B, plan.S, C_out), dtype=torch.float32, device=self._arena.device)
        self._step += 1
        plan.token += 1
        _lib.check(_lib.lib().m2f_plan_forward(plan.handle, self._arena.data_ptr(), self._shadow.data_ptr(), plan.ws.data_ptr(),
                                               None if text is None else text.data_ptr(), None if audio is None else audio.data_ptr(),
                                               mask_u8.data_ptr(), int(dropout_on), self._seed, self._step, logits.data_ptr(), _stream()),
                   "plan_forward")
        self.launches += 1
        return logits

    def _launch_backward(self, plan: _Plan, token: int, dlogits: torch.Tensor):
        if token != plan.token:
            raise RuntimeError("M2FNet backward: the workspace of this (B,S) plan was overwritten by a later forward; "
                               "call backward before the next forward of the same shape")
        dlogits = dlogits.contiguous().float()
        # `_grads_cleared`: FusedAdam.zero_grad() left the p.grad views attached and only promised they are dead
        attached = self._grads_cleared and self._grad_views is not None
        fresh = attached or all(p.grad is None for p in self._plist)
        if self._grad_arena is None:
            self._grad_arena = torch.zeros(self._arena_floats, dtype=torch.float32, device=self._arena.device)
            self._grad_views = None
        if self._grad_views is None:
            self._grad_views = [self._grad_arena[o:o + n].view(p.shape) for p, o, n in self._flat]
        target = self._grad_arena if fresh else torch.empty_like(self._grad_arena)
        tail = self._tail_opt if self._exchange is not None else None
        if self._exchange is not None:
            if not fresh:
                raise RuntimeError("data-parallel M2FNet needs zero_grad(set_to_none=True) (gradient accumulation across "
                                   "backwards is not supported with the in-backward gradient exchange)")
            if tail is not None:
                tail._before_backward()
            # every finished range is all-reduced by kernels inside the backward's graph; nothing to do per range on the host
            _lib.check(_lib.lib().m2f_plan_backward_bucketed(plan.handle, self._arena.data_ptr(), self._shadow.data_ptr(),
                                                             plan.ws.data_ptr(), dlogits.data_ptr(), target.data_ptr(),
                                                             int(self._exchange[3]), None, None, _stream()), "plan_backward (exchange)")
            self.exchange_ranges = int(_lib.lib().m2f_plan_exchange_ranges(plan.handle))
            if tail is not None:
                tail._applied = True  # the parameters (and the bf16 shadow) have been updated by this backward
        elif self._bucket_hook is not None and fresh:
            errors = []

            def _cb(_user, lo, hi, ready):
                try:  # the range is final behind `ready`: collectives launched by the hook queue up after that stream
                    with torch.cuda.stream(torch.cuda.ExternalStream(int(ready))):
                        self._bucket_hook(int(lo), int(hi))
                except BaseException as e:  # an exception cannot cross the C frame
                    errors.append(e)

            cb = _lib.BUCKET_CB(_cb)
            _lib.check(_lib.lib().m2f_plan_backward_bucketed(plan.handle, self._arena.data_ptr(), self._shadow.data_ptr(),
                                                             plan.ws.data_ptr(), dlogits.data_ptr(), target.data_ptr(),
                                                             int(self._bucket_floats), cb, None, _stream()), "plan_backward_bucketed")
            if errors:
                raise errors[0]
        else:
            if self._bucket_hook is not None:
                raise RuntimeError("data-parallel M2FNet needs zero_grad(set_to_none=True) (gradient accumulation across "
                                   "backwards is not supported with bucketed all-reduce)")
            _lib.check(_lib.lib().m2f_plan_backward(plan.handle, self._arena.data_ptr(), self._shadow.data_ptr(), plan.ws.data_ptr(),
                                                    dlogits.data_ptr(), target.data_ptr(), _stream()), "plan_backward")
        self.launches += 1
        if attached:
            self._grads_cleared = False  # the views already point at the arena the kernels just wrote
        elif fresh:
            for p, v in zip(self._plist, self._grad_views):
                if p.requires_grad:
                    p.grad = v
        else:  # gradient accumulation (zero_grad(set_to_none=False) or several backwards per step)
            for p, o, n in self._flat:
                if not p.requires_grad:
                    continue
                g = target[o:o + n].view(p.shape)
                if p.grad is None:
                    p.grad = g.clone()
                else:
                    p.grad.add_(g)
        if self._done_hook is not None:
            self._done_hook()

    def set_grad_hooks(self, on_bucket, bucket_floats: int, on_done=None):
        """Data-parallel plumbing: `on_bucket(lo, hi)` is called (host side, in stream order) each time the gradient
        arena range [lo, hi) is final; `on_done()` once the backward has been enqueued and `p.grad` views are attached."""
        self._bucket_hook, self._bucket_floats, self._done_hook = on_bucket, int(bucket_floats), on_done

    def forward(self, text, audio, mask):
        """text [B,S,Dt], audio [B,S,Da] fp32, mask [B,S] bool (True = padding) -> logits [B,S,C] fp32
        (src/model.py:102-145).  A disabled modality's tensor is ignored, as in the reference."""
        a_on, t_on = bool(self._ccfg.audio.enabled), bool(self._ccfg.text.enabled)
        used = [x for x, on in ((text, t_on), (audio, a_on)) if on] + [mask]
        for x in used:
            if not isinstance(x, torch.Tensor) or not x.is_cuda:
                raise RuntimeError("M2FNet (B200-native) runs on CUDA tensors only: there is no CPU path. "
                                   "Move the model and the batch to a B200 (`.to('cuda')`).")
        if not self._arena.is_cuda:
            raise RuntimeError("M2FNet (B200-native): move the model to the GPU first (`model.to('cuda')`); there is no CPU path")
        B, S = mask.shape
        text = text.contiguous().float() if t_on else None
        audio = audio.contiguous().float() if a_on else None
        if t_on and tuple(text.shape) != (B, S, self._ccfg.text.dim):
            raise RuntimeError(f"text must be [B,S,{self._ccfg.text.dim}], got {tuple(text.shape)}")
        if a_on and tuple(audio.shape) != (B, S, self._ccfg.audio.dim):
            raise RuntimeError(f"audio must be [B,S,{self._ccfg.audio.dim}], got {tuple(audio.shape)}")
        mask_u8 = mask.contiguous()
        mask_u8 = mask_u8.view(torch.uint8) if mask_u8.dtype == torch.bool else (mask_u8 != 0).view(torch.uint8)
        with torch.cuda.device(self._arena.device):
            self._ensure_shadow()
            anchor = self._grad_anchor()
            needs_grad = torch.is_grad_enabled() and anchor is not None
            plan = self._plan(B, S, needs_grad)
            dropout_on = self.training and float(self.config.dropout) > 0.0
            if needs_grad:
                return _M2FStep.apply(self, plan, text, audio, mask_u8, dropout_on, anchor)
            return self._launch_forward(plan, text, audio, mask_u8, dropout_on)

    def _grad_anchor(self):
        """First parameter that requires grad (autograd needs one differentiable input to record the step's node); the
        scan over all parameters is repeated only when the cached one stops requiring grad or nothing was cached."""
        a = self._anchor_box[0]
        if a is not None and a.requires_grad:
            return a
        self._anchor_box[0] = next((p for p in self._plist if p.requires_grad), None)
        return self._anchor_box[0]

    # ------------------------------------------------------------------------------------------ extras
    def flat_parameters(self) -> torch.Tensor:
        """The fp32 parameter arena (every parameter is a view of it) -- for fused Adam and NCCL buckets."""
        return self._arena

    def flat_grads(self) -> Optional[torch.Tensor]:
        return self._grad_arena

    def _attach_exchange(self, plan: "_Plan"):
        pg, scale, ctas, _ = self._exchange
        _lib.check(_lib.lib().m2f_plan_set_exchange(plan.handle, C.byref(pg), float(scale), int(ctas)), "plan_set_exchange")

    def set_exchange(self, peer_group, scale: float = 1.0, ctas: int = 0, bucket_floats: int = 4 << 20):
        """Attach (or with None detach) a peer gradient exchange: the backward all-reduces each finished range of the
        gradient arena itself (m2f_plan_set_exchange).  The gradient arena must already be this rank's copy of the
        symmetric buffer the group describes (`adopt_grad_arena`)."""
        self._exchange = None if peer_group is None else (peer_group, float(scale), int(ctas), int(bucket_floats))
        for pl in self._plans.values():
            if pl.training and pl.handle:
                if self._exchange is None:
                    _lib.check(_lib.lib().m2f_plan_set_exchange(pl.handle, None, 1.0, 0), "plan_set_exchange")
                else:
                    self._attach_exchange(pl)

    def _attach_optimizer(self, plan: "_Plan"):
        o = self._tail_opt
        _lib.check(_lib.lib().m2f_plan_set_optimizer(plan.handle, o._m.data_ptr(), o._v.data_ptr(), o._hp.data_ptr()), "plan_set_optimizer")

    def set_tail_optimizer(self, opt):
        """Attach (or with None detach) a FusedAdam whose update runs behind the backward's gradient exchange
        (m2f_plan_set_optimizer); it only takes effect while an in-graph exchange is attached."""
        self._tail_opt = opt
        for pl in self._plans.values():
            if pl.training and pl.handle:
                if opt is None:
                    _lib.check(_lib.lib().m2f_plan_set_optimizer(pl.handle, None, None, None), "plan_set_optimizer")
                else:
                    self._attach_optimizer(pl)

    def adopt_grad_arena(self, buf: Optional[torch.Tensor]):
        """Make the backward write its gradients into `buf` (fp32, arena_floats elements, on the model's device) -- the
        peer gradient exchange places the arena in a symmetric allocation this way.  None reverts to a private arena."""
        if buf is not None:
            if buf.dtype != torch.float32 or buf.numel() != self._arena_floats or buf.device != self._arena.device or not buf.is_contiguous():
                raise ValueError("adopt_grad_arena: need a contiguous fp32 tensor of arena size on the model's device")
        for p in self._plist:
            p.grad = None
        self._grad_arena, self._grad_views, self._grads_cleared = buf, None, False

    def __del__(self):
        try:
            for pl in self._plans.values():
                pl.destroy()
            if getattr(self, "_handle", None):
                _lib.lib().m2f_model_destroy(self._handle)
                self._handle = None
        except Exception:
            pass


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, weight, eps):
        from . import ops
        B, S, Cc = logits.shape
        lg = logits.reshape(B * S, Cc)
        dl = torch.empty_like(lg)
        out = ops.ce_fwd_bwd(lg, target.reshape(-1).contiguous(), weight, eps=eps, grad_scale=1.0, dlogits=dl)
        ctx.save_for_backward(dl)
        ctx.shape = (B, S, Cc)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return (dl * g).view(ctx.shape), None, None, None


class FusedCrossEntropyLoss(nn.Module):
    """CrossEntropyLoss(weight?, ignore_index=-1, label_smoothing=eps) as src/train.py:48-50 builds it, computed by
    the K4 kernel (loss and dlogits in one pass).  Called like the reference's criterion: `criterion(outputs.permute(0,2,1), emotion)`."""

    def __init__(self, weight: Optional[torch.Tensor] = None, ignore_index: int = -1, label_smoothing: float = 0.1):
        super().__init__()
        if ignore_index != -1:
            raise ValueError("FusedCrossEntropyLoss supports ignore_index=-1 only (src/train.py:48-50)")
        self.register_buffer("weight", weight.float() if weight is not None else None)
        self.eps = float(label_smoothing)

    def forward(self, outputs_bcs: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if not outputs_bcs.is_cuda:
            raise RuntimeError("FusedCrossEntropyLoss runs on CUDA tensors only")
        logits = outputs_bcs.permute(0, 2, 1)  # back to [B,S,C]; a no-copy view of the model's logits
        if not logits.is_contiguous():
            logits = logits.contiguous()
        return _CEFn.apply(logits.float(), target, self.weight, self.eps)
