"""K8 — LayerNorm / RMSNorm (csrc/kernels/norm.cu) as autograd functions.

``norm(x, gamma, beta=None, residual=None, rms=False)`` returns ``y`` or, when a
residual is given, ``(y, h)`` with ``h = x + residual`` (the fused
add-then-normalise of a pre-LN transformer block).
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import c_float, c_int, ptr, stream_ptr


def norm_fwd_raw(x2, residual2, gamma, beta, eps, rms):
    rows, H = x2.shape
    y = torch.empty_like(x2)
    h = torch.empty_like(x2) if residual2 is not None else None
    mean = None if rms else torch.empty(rows, dtype=torch.float32, device=x2.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x2.device)
    _lib.call("b2f_norm_fwd", ptr(x2), ptr(residual2), ptr(gamma), ptr(beta), ptr(y), ptr(h),
              ptr(mean), ptr(rstd), c_int(rows), c_int(H), c_float(eps), c_int(int(rms)),
              stream_ptr())
    _lib.debug_sync()
    return y, h, mean, rstd


def norm_bwd_raw(dy2, x2, gamma, mean, rstd, dres2, rms, dgamma=None, dbeta=None,
                 accumulate=False, scratch=None):
    rows, H = x2.shape
    dx = torch.empty_like(x2)
    nparts = _lib.call_int("b2f_norm_bwd_num_partials", c_int(rows))
    if scratch is None:
        scratch = torch.empty(nparts * 2 * H, dtype=torch.float32, device=x2.device)
    if dgamma is None:
        dgamma = torch.empty(H, dtype=torch.float32, device=x2.device)
        accumulate = False
    if dbeta is None and not rms:
        dbeta = torch.empty(H, dtype=torch.float32, device=x2.device)
    _lib.call("b2f_norm_bwd", ptr(dy2), ptr(x2), ptr(gamma), ptr(mean), ptr(rstd), ptr(dres2),
              ptr(dx), ptr(dgamma), ptr(dbeta), ptr(scratch), c_int(rows), c_int(H),
              c_int(int(rms)), c_int(int(accumulate)), stream_ptr())
    _lib.debug_sync()
    return dx, dgamma, dbeta


class _Norm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, eps, rms):
        _lib.require_cuda(x)
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        r2 = residual.reshape(-1, x.shape[-1]).contiguous() if residual is not None else None
        y, h, mean, rstd = norm_fwd_raw(x2, r2, gamma, beta, eps, rms)
        inp = h if h is not None else x2
        ctx.save_for_backward(inp, gamma, mean if mean is not None else rstd, rstd)
        ctx.rms, ctx.has_res, ctx.has_beta = rms, residual is not None, beta is not None
        ctx.shape = x.shape
        ctx.beta_ref = beta
        if residual is not None:
            return y.reshape(x.shape), h.reshape(x.shape)
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy, dh=None):
        inp, gamma, mean, rstd = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres2 = None
        if ctx.has_res and dh is not None:
            dres2 = dh.reshape(-1, dh.shape[-1])
            if not dres2.is_contiguous():
                dres2 = dres2.contiguous()
        mg = getattr(gamma, "main_grad", None)
        mb = getattr(ctx.beta_ref, "main_grad", None) if ctx.beta_ref is not None else None
        if mg is not None and not ctx.rms and mb is None:
            mb = torch.zeros_like(mg)
        dx, dgamma, dbeta = norm_bwd_raw(dy2, inp, gamma, None if ctx.rms else mean, rstd, dres2,
                                         ctx.rms, dgamma=mg, dbeta=mb, accumulate=mg is not None)
        dx = dx.reshape(ctx.shape)
        if mg is not None:
            for prm in (gamma, ctx.beta_ref):
                cb = getattr(prm, "grad_ready", None) if prm is not None else None
                if cb is not None and getattr(prm, "main_grad", None) is not None:
                    cb()
        dg = None if mg is not None else dgamma.to(gamma.dtype)
        has_mb = ctx.beta_ref is not None and getattr(ctx.beta_ref, "main_grad", None) is not None
        db = dbeta.to(gamma.dtype) if (ctx.has_beta and dbeta is not None and not has_mb) else None
        return dx, dg, db, (dx if ctx.has_res else None), None, None


def norm(x, gamma, beta=None, residual=None, eps: float = 1e-5, rms: bool = False):
    if not x.is_cuda:
        from . import ref
        return ref.norm(x, gamma, beta, residual, eps, rms)
    return _Norm.apply(x, gamma, beta, residual, eps, rms)


def rmsnorm(x, gamma, residual=None, eps: float = 1e-5):
    return norm(x, gamma, None, residual, eps, True)


def layernorm(x, gamma, beta, residual=None, eps: float = 1e-5):
    return norm(x, gamma, beta, residual, eps, False)
