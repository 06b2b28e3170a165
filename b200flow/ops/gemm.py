"""K0 — tcgen05 GEMM and the linear-layer autograd functions built on it.

`gemm()` is the thin wrapper over ``b2f_gemm_bf16`` (csrc/kernels/gemm.cu).
`linear()` is y = x Wᵀ (+ bias) with a backward made of two more calls of the same
kernel: dgrad (B operand MN-major) and wgrad (both operands MN-major) — no
transposes are materialised.  When the weight carries a ``main_grad`` fp32 buffer
the wgrad GEMM accumulates straight into it (beta = 1 epilogue), which is what
apex's fused_weight_gradient_mlp_cuda does for Megatron.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import c_float, c_int, c_ll, ptr, stream_ptr

import os

EPI_GELU, EPI_OUT_F32, EPI_ACCUM, EPI_RELU, EPI_BIAS_ROW, EPI_DGELU = 1, 2, 4, 8, 16, 32
# "2cta": CTA-pair kernel (cta_group::2) for tiles that fill it; "1cta": always the 128x256 kernel
GEMM_IMPL = os.environ.get("B2F_GEMM", "2cta")


def _entry(M: int, N: int) -> str:
    if GEMM_IMPL == "2cta" and M >= 256 and N >= 256:
        return "b2f_gemm_bf16_2cta"
    return "b2f_gemm_bf16"


def _ld(t: torch.Tensor) -> int:
    assert t.stride(-1) == 1, "innermost dimension must be contiguous"
    return t.stride(-2)


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
         out: torch.Tensor | None = None, out_dtype: torch.dtype | None = None,
         bias: torch.Tensor | None = None, residual: torch.Tensor | None = None,
         aux: torch.Tensor | None = None, gelu: bool = False, relu: bool = False,
         accumulate: bool = False, alpha: float = 1.0, dgelu: bool = False) -> torch.Tensor:
    """C = epilogue(alpha * op(A) @ op(B)ᵀ-style product), see include/b200flow.h.

    a: [.., M, K] (or [.., K, M] if a_mn_major);  b: [.., N, K] (or [.., K, N] if b_mn_major).
    """
    _lib.require_cuda(a, b)
    assert a.dtype == _lib.HALF and b.dtype == _lib.HALF
    assert a.dim() == b.dim() and a.dim() in (2, 3)
    batch = a.shape[0] if a.dim() == 3 else 1
    if a_mn_major:
        K, M = a.shape[-2], a.shape[-1]
    else:
        M, K = a.shape[-2], a.shape[-1]
    if b_mn_major:
        Kb, N = b.shape[-2], b.shape[-1]
    else:
        N, Kb = b.shape[-2], b.shape[-1]
    assert K == Kb, f"contraction mismatch {K} vs {Kb}"
    if out_dtype is None:
        out_dtype = _lib.HALF
    if out is None:
        shape = (batch, M, N) if a.dim() == 3 else (M, N)
        out = torch.empty(shape, dtype=out_dtype, device=a.device)
        assert not accumulate, "accumulate needs an explicit fp32 `out`"
    flags = 0
    if out.dtype == torch.float32:
        flags |= EPI_OUT_F32
    else:
        assert out.dtype == _lib.HALF
    if accumulate:
        flags |= EPI_ACCUM
    if gelu:
        flags |= EPI_GELU
    if relu:
        flags |= EPI_RELU
    if dgelu:   # `residual` is a saved pre-activation z: C = (A B) * gelu'(z)
        assert residual is not None
        flags |= EPI_DGELU
    sa = a.stride(0) if a.dim() == 3 else 0
    sb = b.stride(0) if b.dim() == 3 else 0
    sc = out.stride(0) if out.dim() == 3 else 0
    sr = residual.stride(0) if (residual is not None and residual.dim() == 3) else 0
    sx = aux.stride(0) if (aux is not None and aux.dim() == 3) else 0
    _lib.call(
        _entry(M, N), ptr(a), ptr(b), ptr(out), c_int(M), c_int(N), c_int(K),
        c_ll(_ld(a)), c_ll(_ld(b)), c_ll(_ld(out)), c_int(int(a_mn_major)), c_int(int(b_mn_major)),
        c_int(batch), c_ll(sa), c_ll(sb), c_ll(sc), ptr(bias), ptr(residual),
        c_ll(_ld(residual) if residual is not None else 0), c_ll(sr), ptr(aux),
        c_ll(_ld(aux) if aux is not None else 0), c_ll(sx), c_float(alpha), c_int(flags),
        stream_ptr())
    _lib.debug_sync()
    return out


# ---- deferred weight gradients (pipeline bubble filling) --------------------------------------------------------
# A backward pass is two independent halves per linear: dgrad (needed by the previous layer / pipeline stage right
# away) and wgrad (needed only by the optimizer).  Inside `defer_wgrad(queue)` every wgrad GEMM whose target is a
# `main_grad` buffer is queued instead of launched; `run_deferred(queue)` launches them later - the pipeline schedule
# runs them where a stage would otherwise idle (parallel/pipeline.py, "W" instructions).
DEFER_CAPABLE = False            # engines that split backward set this, so CPU linears take the same autograd node
_DEFERRED: list | None = None


class defer_wgrad:
    def __init__(self, queue: list):
        self.queue = queue

    def __enter__(self):
        global _DEFERRED
        self._prev, _DEFERRED = _DEFERRED, self.queue
        return self.queue

    def __exit__(self, *exc):
        global _DEFERRED
        _DEFERRED = self._prev


def run_deferred(queue: list) -> None:
    global _DEFERRED
    prev, _DEFERRED = _DEFERRED, None
    try:
        with torch.no_grad():
            for x2, dy2, weight in queue:
                _wgrad(x2, dy2, weight)
    finally:
        _DEFERRED = prev
        queue.clear()


def _wgrad(x2: torch.Tensor, dy2: torch.Tensor, weight: torch.Tensor):
    """dW[N, K] = dyᵀ x, accumulated into weight.main_grad when present."""
    main_grad = getattr(weight, "main_grad", None)
    if main_grad is not None:
        if _DEFERRED is not None:
            _DEFERRED.append((x2, dy2, weight))
            return None
        if dy2.is_cuda:
            gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=main_grad, accumulate=True)
        else:
            main_grad.add_((dy2.t().to(main_grad.dtype) @ x2.to(main_grad.dtype)))
        cb = getattr(weight, "grad_ready", None)   # DDP bucket bookkeeping
        if cb is not None:
            cb()
        return None
    if not dy2.is_cuda:
        return dy2.t() @ x2
    return gemm(dy2, x2, a_mn_major=True, b_mn_major=True)


class _LinearCPU(torch.autograd.Function):
    """The CPU (oracle) linear as ONE node with the same dgrad / wgrad split as the GPU node, so that schedules which
    defer weight gradients are testable over gloo."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual):
        x2 = x.reshape(-1, x.shape[-1])
        y = torch.nn.functional.linear(x2, weight, bias)
        if residual is not None:
            y = y + residual.reshape(-1, weight.shape[0])
        ctx.save_for_backward(x2, weight)
        ctx.has_bias, ctx.has_res, ctx.x_shape = bias is not None, residual is not None, x.shape
        # 2-D input: hand out y itself, not a same-shape view of it - the vocab-parallel loss turns the logits into their
        # gradient in place, which autograd refuses on a view created inside a custom Function
        return y if x.dim() == 2 else y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        dx = (dy2 @ weight).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = _wgrad(x2, dy2, weight)
        return dx, dw, (dy2.sum(0) if ctx.has_bias else None), (dy if ctx.has_res else None)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, residual):
        x2 = x.reshape(-1, x.shape[-1])
        r2 = residual.reshape(-1, weight.shape[0]) if residual is not None else None
        y = gemm(x2, weight, bias=bias, residual=r2)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.has_res = residual is not None
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = gemm(dy2, weight, b_mn_major=True).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = _wgrad(x2, dy2, weight)
        db = dy2.sum(0) if ctx.has_bias else None
        return dx, dw, db, (dy if ctx.has_res else None)


def linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None,
           residual: torch.Tensor | None = None) -> torch.Tensor:
    """y = x @ weightᵀ (+ bias) (+ residual) on the tcgen05 GEMM."""
    if not x.is_cuda:
        if DEFER_CAPABLE and getattr(weight, "main_grad", None) is not None:
            return _LinearCPU.apply(x, weight, bias, residual)
        from . import ref
        return ref.linear(x, weight, bias, residual)
    return _Linear.apply(x, weight, bias, residual)


class _LinearGelu(torch.autograd.Function):
    """y = gelu(x Wᵀ + b) with the pre-activation saved by the same GEMM epilogue."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = x.reshape(-1, x.shape[-1])
        pre = torch.empty((x2.shape[0], weight.shape[0]), dtype=_lib.HALF, device=x.device)
        y = gemm(x2, weight, bias=bias, aux=pre, gelu=True)
        ctx.save_for_backward(x2, weight, pre)
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        from .elementwise import gelu_bwd_raw

        x2, weight, pre = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dpre = gelu_bwd_raw(pre, None, dy2)
        dx = gemm(dpre, weight, b_mn_major=True).reshape(ctx.x_shape)
        dw = _wgrad(x2, dpre, weight)
        return dx, dw, dpre.sum(0)


class _MlpGelu(torch.autograd.Function):
    """y = gelu(x W1ᵀ + b1) W2ᵀ + b2 as one autograd node, so the backward can fold gelu' into the
    epilogue of the dgrad GEMM through W2 (no separate activation-gradient pass over [tokens, ffn])."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2):
        x2 = x.reshape(-1, x.shape[-1])
        pre = torch.empty((x2.shape[0], w1.shape[0]), dtype=_lib.HALF, device=x.device)
        h = gemm(x2, w1, bias=b1, aux=pre, gelu=True)
        y = gemm(h, w2, bias=b2)
        ctx.save_for_backward(x2, w1, w2, pre, h)
        ctx.x_shape = x.shape
        ctx.has_b = (b1 is not None, b2 is not None)
        return y.reshape(*x.shape[:-1], w2.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w1, w2, pre, h = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dpre = gemm(dy2, w2, b_mn_major=True, residual=pre, dgelu=True)
        dw2 = _wgrad(h, dy2, w2)
        db2 = dy2.sum(0) if ctx.has_b[1] else None
        dx = gemm(dpre, w1, b_mn_major=True).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw1 = _wgrad(x2, dpre, w1)
        db1 = dpre.sum(0) if ctx.has_b[0] else None
        return dx, dw1, db1, dw2, db2


def mlp_gelu(x, w1, b1, w2, b2):
    """Transformer FFN with GELU: two tcgen05 GEMMs forward, four backward, no elementwise pass."""
    if not x.is_cuda:
        from . import ref
        return ref.linear(ref.linear_gelu(x, w1, b1), w2, b2, None)
    return _MlpGelu.apply(x, w1, b1, w2, b2)


def linear_gelu(x, weight, bias):
    if not x.is_cuda:
        from . import ref
        return ref.linear_gelu(x, weight, bias)
    return _LinearGelu.apply(x, weight, bias)
