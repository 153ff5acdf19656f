"""Fused Adam over the model's flat parameter arena: one kernel per step instead of
torch.optim.Adam's foreach launches over 186 tensors.  Same update rule as the reference's optimizer,
`torch.optim.Adam(model.parameters(), lr, weight_decay=wd)` (src/train.py:56,231): L2 decay folded into the gradient,
bias-corrected moments, eps 1e-8.  `state_dict()` has torch.optim.Adam's layout (step / exp_avg / exp_avg_sq per
parameter), so checkpoints interchange with the reference's `optimizer_state_dict` (src/train.py:163-168)."""
from __future__ import annotations

import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, grad_scale: float = 1.0,
                 overlap_exchange: bool = True):
        """`overlap_exchange` (data parallel with the in-graph peer exchange only; no effect otherwise): the update of every
        exchanged gradient range is enqueued by `loss.backward()` itself, behind the backward's compute and gated by the
        exchange of the NEXT range, so that it overlaps the tail of the exchange instead of waiting for all of it
        (m2f_plan_set_optimizer); `step()` then only advances the step count.  Bit-identical results.  What it asks of the
        caller: every backward() is followed by one step() before the next forward (the reference's loop, src/train.py:228-231);
        a second backward() without step() raises.  The learning rate of `param_groups[0]` is read at each backward."""
        self.model = getattr(model, "module", model)
        params = [p for p, _, _ in self.model._flat]
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.grad_scale = float(grad_scale)
        self._t = 0
        self._applied = False
        self._bind()
        self.overlap_exchange = bool(overlap_exchange)
        if self.overlap_exchange:
            self._hp = torch.zeros(8, dtype=torch.float32, device=self.model.flat_parameters().device)  # M2F_ADAM_HYPER_FLOATS
            self.model.set_tail_optimizer(self)

    def _before_backward(self):
        if self._applied:
            raise RuntimeError("FusedAdam(overlap_exchange=True): backward() ran twice without step() in between -- the first "
                               "backward has already applied its update")
        grp = self.param_groups[0]
        _lib.check(_lib.lib().m2f_adam_hyper(self._hp.data_ptr(), float(grp["lr"]), float(grp["betas"][0]), float(grp["betas"][1]),
                                             float(grp["eps"]), float(grp["weight_decay"]), self._t + 1, self.grad_scale,
                                             torch.cuda.current_stream().cuda_stream), "adam_hyper")

    def detach_from_backward(self):
        """Back to updating in step() only."""
        self.model.set_tail_optimizer(None)
        self.overlap_exchange = False

    def _bind(self):
        arena = self.model.flat_parameters()
        if not arena.is_cuda:
            raise RuntimeError("FusedAdam needs the model on the GPU (there is no CPU path)")
        self._m = torch.zeros_like(arena)
        self._v = torch.zeros_like(arena)
        for p, o, n in self.model._flat:
            self.state[p] = {"step": torch.tensor(float(self._t)), "exp_avg": self._m[o:o + n].view(p.shape),
                             "exp_avg_sq": self._v[o:o + n].view(p.shape)}

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        g = self.model.flat_grads()
        if g is None:
            raise RuntimeError("FusedAdam.step(): no gradients (call backward first)")
        if self._applied:  # the backward has already applied this update behind its gradient exchange
            self._applied = False
            self._t += 1
            return loss
        grp = self.param_groups[0]
        self._t += 1
        arena = self.model.flat_parameters()
        mdl = self.model
        hp = (float(grp["lr"]), float(grp["betas"][0]), float(grp["betas"][1]), float(grp["eps"]), float(grp["weight_decay"]), self._t,
              self.grad_scale, torch.cuda.current_stream().cuda_stream)
        if mdl._shadow is not None and mdl._shadow_version == mdl._param_version() and mdl._views_intact():
            # shadow is current: update and re-split the weights in one pass; it stays current
            _lib.check(_lib.lib().m2f_model_adam_step(mdl._handle, arena.data_ptr(), g.data_ptr(), self._m.data_ptr(), self._v.data_ptr(),
                                                      mdl._shadow.data_ptr(), *hp), "model_adam_step")
        else:
            _lib.check(_lib.lib().m2f_adam_step(arena.data_ptr(), g.data_ptr(), self._m.data_ptr(), self._v.data_ptr(), arena.numel(), *hp),
                       "adam_step")
            mdl._shadow_version = None  # the kernel wrote through raw pointers: the bf16 weight shadow is stale
        return loss

    def zero_grad(self, set_to_none: bool = True):
        """With `set_to_none=True` (the default, and what the reference's loop does, src/train.py:225) the gradient views
        stay attached to the arena and are only declared dead: the next backward overwrites the arena in place, so
        nothing has to be detached and re-attached for 186 parameters every step.  Until that backward `p.grad` still
        shows the previous step's values."""
        m = self.model
        if set_to_none and m._grad_views is not None and all(p.grad is v for p, v in zip(m._plist[:4], m._grad_views[:4])):
            m._grads_cleared = True
            return
        m._grads_cleared = False
        super().zero_grad(set_to_none=set_to_none)

    def state_dict(self):
        for st in self.state.values():
            st["step"] = torch.tensor(float(self._t))
        return super().state_dict()

    def load_state_dict(self, sd):
        super().load_state_dict(sd)
        first = True
        for p, o, n in self.model._flat:
            st = self.state[p]
            if first:
                self._t = int(float(st["step"]))
                first = False
            self._m[o:o + n].view(p.shape).copy_(st["exp_avg"])
            self._v[o:o + n].view(p.shape).copy_(st["exp_avg_sq"])
            st["exp_avg"], st["exp_avg_sq"] = self._m[o:o + n].view(p.shape), self._v[o:o + n].view(p.shape)
