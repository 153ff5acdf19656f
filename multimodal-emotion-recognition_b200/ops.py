"""Op-level Python wrappers over the C ABI (used by the parity tests and the host-side model).

Every wrapper takes CUDA torch tensors, extracts raw device pointers / strides and the current CUDA
stream, and calls straight into libm2f_b200.so.  Tensors on the CPU are rejected: there is no fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import Epilogue, KSeg, Planes


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("m2f_b200 ops need CUDA tensors: this package has no CPU path (SURVEY.md 8b)")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class PlanesTensor:
    """A [planes, rows, ld] bf16 tensor viewed as a logical [rows, cols] matrix (cols <= ld)."""

    def __init__(self, data: torch.Tensor, cols: int, col0: int = 0):
        assert data.dtype == torch.bfloat16 and data.dim() == 3 and data.is_contiguous()
        self.data, self.cols, self.col0 = data, cols, col0

    @classmethod
    def empty(cls, rows: int, cols: int, planes: int, device, ld: Optional[int] = None, zero: bool = False):
        ld = ld or (cols + 7) // 8 * 8
        f = torch.zeros if zero else torch.empty
        return cls(f((planes, rows, ld), dtype=torch.bfloat16, device=device), cols)

    @classmethod
    def from_float(cls, x: torch.Tensor, planes: int, ld: Optional[int] = None):
        """fp32 [rows, cols] -> planes through the CUDA kernel m2f_split_planes."""
        _need_cuda(x)
        x = x.contiguous().float()
        out = cls.empty(x.shape[0], x.shape[1], planes, x.device, ld, zero=True)
        _lib.check(_lib.lib().m2f_split_planes(x.data_ptr(), x.stride(0), C.byref(out.c()), _stream()), "split_planes")
        return out

    def cview(self, col0: int, cols: int) -> "PlanesTensor":
        return PlanesTensor(self.data, cols, self.col0 + col0)

    @property
    def rows(self):
        return self.data.shape[1]

    @property
    def planes(self):
        return self.data.shape[0]

    def c(self) -> Planes:
        d = self.data
        return Planes(d.data_ptr() + 2 * self.col0, d.shape[1], self.cols, d.stride(1), d.stride(0), d.shape[0], 0)

    def float(self) -> torch.Tensor:
        return self.data[:, :, self.col0:self.col0 + self.cols].float().sum(dim=0)


def gemm(A: PlanesTensor, B: PlanesTensor, M: int, N: int, *, a_mn: bool = False, b_mn: bool = False,
         segs: Optional[Sequence[Tuple[int, int, int]]] = None, bias: Optional[torch.Tensor] = None, relu: bool = False,
         gate: Optional[PlanesTensor] = None, gate_scale: float = 1.0, drop_p: float = 0.0, seed: int = 0, step: int = 0,
         site: int = 0, drop_ld: Optional[int] = None, resid: Optional[torch.Tensor] = None, out_f32: Optional[torch.Tensor] = None,
         accumulate: bool = False, out_planes: Optional[PlanesTensor] = None, colsum: Optional[torch.Tensor] = None):
    """D = A . B^T (+ fused epilogue), see m2f_gemm in include/m2f_b200.h."""
    _need_cuda(A.data, B.data, bias, resid, out_f32)
    if segs is None:
        K = A.rows if a_mn else A.cols
        segs = [(0, 0, K)]
    cs = (KSeg * len(segs))(*[KSeg(*s) for s in segs])
    e = Epilogue()
    e.bias = _ptr(bias)
    e.relu = int(relu)
    e.accumulate = int(accumulate)
    if gate is not None:
        e.gate = gate.data.data_ptr() + 2 * gate.col0
        e.gate_ld = gate.data.stride(1)
    e.gate_scale = gate_scale
    e.drop_p = drop_p
    e.seed, e.step, e.site = seed, step, site
    e.drop_ld = N if drop_ld is None else drop_ld
    if resid is not None:
        e.resid, e.resid_ld = resid.data_ptr(), resid.stride(0)
    if out_f32 is not None:
        e.out_f32, e.out_f32_ld = out_f32.data_ptr(), out_f32.stride(0)
    if out_planes is not None:
        e.out_planes = out_planes.c()
    e.colsum = _ptr(colsum)
    ca, cb = A.c(), B.c()
    _lib.check(_lib.lib().m2f_gemm(C.byref(ca), int(a_mn), C.byref(cb), int(b_mn), cs, len(segs), M, N, C.byref(e), _stream()), "gemm")


def relu_planes(src: PlanesTensor, dst: PlanesTensor):
    cs, cd = src.c(), dst.c()
    _lib.check(_lib.lib().m2f_relu_planes(C.byref(cs), C.byref(cd), _stream()), "relu_planes")


def colsum_planes(src: PlanesTensor, out: torch.Tensor):
    _need_cuda(out)
    cs = src.c()
    _lib.check(_lib.lib().m2f_colsum_planes(C.byref(cs), out.data_ptr(), _stream()), "colsum_planes")


def ln_fwd(x, gamma, beta, *, skip=None, y=None, y_planes: Optional[PlanesTensor] = None, drop_p=0.0, seed=0, step=0, site=0, stats=None):
    _need_cuda(x, gamma, beta, skip, y, stats)
    M, D = x.shape
    yp = y_planes.c() if y_planes is not None else Planes()
    _lib.check(_lib.lib().m2f_ln_fwd(x.data_ptr(), x.stride(0), gamma.data_ptr(), beta.data_ptr(), _ptr(skip),
                                     skip.stride(0) if skip is not None else 0, M, D, _ptr(y), y.stride(0) if y is not None else 0,
                                     C.byref(yp), drop_p, seed, step, site, _ptr(stats), _stream()), "ln_fwd")


def ln_bwd(g, x, gamma, stats, *, dx=None, dx_planes: Optional[PlanesTensor] = None, drop_p=0.0, seed=0, step=0, site=0,
           dgamma=None, dbeta=None, dbias=None):
    _need_cuda(g, x, gamma, stats, dx, dgamma, dbeta, dbias)
    M, D = x.shape
    dp = dx_planes.c() if dx_planes is not None else Planes()
    _lib.check(_lib.lib().m2f_ln_bwd(g.data_ptr(), g.stride(0), x.data_ptr(), x.stride(0), gamma.data_ptr(), stats.data_ptr(), M, D,
                                     _ptr(dx), dx.stride(0) if dx is not None else 0, C.byref(dp), drop_p, seed, step, site,
                                     _ptr(dgamma), _ptr(dbeta), _ptr(dbias), _stream()), "ln_bwd")


def attn_fwd(q: PlanesTensor, k: PlanesTensor, v: PlanesTensor, key_pad: torch.Tensor, B: int, S: int, H: int, o: PlanesTensor,
             drop_p=0.0, seed=0, step=0, site=0):
    _need_cuda(key_pad)
    assert key_pad.dtype == torch.uint8 and key_pad.is_contiguous()
    cq, ck, cv, co = q.c(), k.c(), v.c(), o.c()
    _lib.check(_lib.lib().m2f_attn_fwd(C.byref(cq), C.byref(ck), C.byref(cv), key_pad.data_ptr(), B, S, H, drop_p, seed, step, site,
                                       C.byref(co), _stream()), "attn_fwd")


def attn_bwd(q, k, v, d_o, key_pad, B, S, H, dq, dk, dv, drop_p=0.0, seed=0, step=0, site=0, dbq=None, dbk=None, dbv=None):
    _need_cuda(key_pad, dbq, dbk, dbv)
    cq, ck, cv, cdo, cdq, cdk, cdv = q.c(), k.c(), v.c(), d_o.c(), dq.c(), dk.c(), dv.c()
    _lib.check(_lib.lib().m2f_attn_bwd(C.byref(cq), C.byref(ck), C.byref(cv), C.byref(cdo), key_pad.data_ptr(), B, S, H, drop_p, seed,
                                       step, site, C.byref(cdq), C.byref(cdk), C.byref(cdv), _ptr(dbq), _ptr(dbk), _ptr(dbv), _stream()), "attn_bwd")


def ce_fwd_bwd(logits: torch.Tensor, target: torch.Tensor, class_w: Optional[torch.Tensor] = None, eps: float = 0.1,
               grad_scale: float = 1.0, dlogits: Optional[torch.Tensor] = None, dl_planes: Optional[PlanesTensor] = None):
    """Returns a 4-float device tensor: [loss, sum of weights, nll sum, smoothing sum]."""
    _need_cuda(logits, target, class_w, dlogits)
    M, Cc = logits.shape
    assert target.dtype == torch.int64 and target.is_contiguous()
    out = torch.empty(4, dtype=torch.float32, device=logits.device)
    dp = dl_planes.c() if dl_planes is not None else Planes()
    _lib.check(_lib.lib().m2f_ce_fwd_bwd(logits.data_ptr(), logits.stride(0), target.data_ptr(), _ptr(class_w), M, Cc, eps, grad_scale,
                                         out.data_ptr(), _ptr(dlogits), dlogits.stride(0) if dlogits is not None else 0, C.byref(dp),
                                         _stream()), "ce_fwd_bwd")
    return out


def ce_fwd_bwd_static(logits: torch.Tensor, target: torch.Tensor, out: torch.Tensor, class_w: Optional[torch.Tensor] = None,
                      eps: float = 0.1, grad_scale: float = 1.0, dlogits: Optional[torch.Tensor] = None,
                      dl_planes: Optional[PlanesTensor] = None) -> torch.Tensor:
    """m2f_ce_fwd_bwd_static: the allocation-free cross entropy.  `out` (2 floats, device) is the caller's: [loss, sum of
    weights]; this wrapper allocates nothing either.  M <= M2F_CE_STATIC_MAX_ROWS; calls sharing a GPU must not overlap."""
    _need_cuda(logits, target, out, class_w, dlogits)
    M, Cc = logits.shape
    assert target.dtype == torch.int64 and target.is_contiguous()
    assert out.dtype == torch.float32 and out.numel() >= 2 and out.is_contiguous()
    dp = dl_planes.c() if dl_planes is not None else Planes()
    _lib.check(_lib.lib().m2f_ce_fwd_bwd_static(logits.data_ptr(), logits.stride(0), target.data_ptr(), _ptr(class_w), M, Cc, eps,
                                                grad_scale, out.data_ptr(), _ptr(dlogits),
                                                dlogits.stride(0) if dlogits is not None else 0, C.byref(dp), _stream()),
               "ce_fwd_bwd_static")
    return out


def dropout_mask(p: float, seed: int, step: int, site: int, n: int, device) -> torch.Tensor:
    m = torch.empty(n, dtype=torch.uint8, device=device)
    _need_cuda(m)
    _lib.check(_lib.lib().m2f_dropout_mask(p, seed, step, site, n, m.data_ptr(), _stream()), "dropout_mask")
    return m
