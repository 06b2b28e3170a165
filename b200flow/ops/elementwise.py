"""K10 + RoPE + cross-entropy wrappers (csrc/kernels/elementwise.cu)."""
from __future__ import annotations

import torch

from . import _lib
from ._lib import c_float, c_int, c_ll, c_ull, ptr, stream_ptr


def colsum(x: torch.Tensor) -> torch.Tensor:
    """x.sum(0) of a half-precision [rows, C] matrix with fp32 accumulation: one pass (`b2f_colsum_partial`) plus a small
    deterministic reduction of the per-block partial rows.  Used for convolution bias gradients over NHWC memory, where
    torch's reduction is 2-3x slower; for row-major linear-layer gradients torch's own sum is as fast (measured 49 vs
    54 us at 32768 x 1024), so those keep `.sum(0)`.  Falls back to torch for other dtypes / layouts / CPU."""
    if not (x.is_cuda and x.dim() == 2 and x.dtype == _lib.HALF and x.is_contiguous() and x.shape[1] % 8 == 0
            and x.shape[0] > 0):
        return x.sum(0)
    rows, C = x.shape
    nb = _lib.cuda_lib().b2f_colsum_num_partials(c_ll(rows), c_int(C))
    partial = torch.empty((nb, C), dtype=torch.float32, device=x.device)
    _lib.call("b2f_colsum_partial", ptr(x), ptr(partial), c_ll(rows), c_int(C), stream_ptr())
    return partial.sum(0).to(x.dtype)


# ------------------------------------------------------------------ SwiGLU
class _SwiGLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        _lib.require_cuda(gu)
        g2 = gu.reshape(-1, gu.shape[-1])
        if not g2.is_contiguous():
            g2 = g2.contiguous()
        F = g2.shape[1] // 2
        out = torch.empty((g2.shape[0], F), dtype=gu.dtype, device=gu.device)
        _lib.call("b2f_swiglu_fwd", ptr(g2), ptr(out), c_ll(g2.shape[0]), c_int(F), stream_ptr())
        ctx.save_for_backward(g2)
        ctx.shape = gu.shape
        return out.reshape(*gu.shape[:-1], F)

    @staticmethod
    def backward(ctx, dout):
        (g2,) = ctx.saved_tensors
        F = g2.shape[1] // 2
        d2 = dout.reshape(-1, F)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        dgu = torch.empty_like(g2)
        _lib.call("b2f_swiglu_bwd", ptr(g2), ptr(d2), ptr(dgu), c_ll(g2.shape[0]), c_int(F),
                  stream_ptr())
        return dgu.reshape(ctx.shape)


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    """out = silu(gu[..., :F]) * gu[..., F:]."""
    if not gu.is_cuda:
        from . import ref
        return ref.swiglu(gu)
    return _SwiGLU.apply(gu)


# -------------------------------------------------------------------- GELU
def gelu_fwd_raw(x2, bias):
    y = torch.empty_like(x2)
    _lib.call("b2f_gelu_fwd", ptr(x2), ptr(bias), ptr(y), c_ll(x2.shape[0]), c_int(x2.shape[1]),
              stream_ptr())
    return y


def gelu_bwd_raw(x2, bias, dy2):
    dx = torch.empty_like(x2)
    _lib.call("b2f_gelu_bwd", ptr(x2), ptr(bias), ptr(dy2), ptr(dx), c_ll(x2.shape[0]),
              c_int(x2.shape[1]), stream_ptr())
    return dx


class _BiasGelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        ctx.save_for_backward(x2, bias)
        ctx.shape = x.shape
        return gelu_fwd_raw(x2, bias).reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, bias = ctx.saved_tensors
        dx = gelu_bwd_raw(x2, bias, dy.reshape(-1, dy.shape[-1]).contiguous())
        return dx.reshape(ctx.shape), (dx.sum(0) if bias is not None else None)


def bias_gelu(x, bias=None):
    if not x.is_cuda:
        from . import ref
        return ref.bias_gelu(x, bias)
    return _BiasGelu.apply(x, bias)


# ------------------------------------------------------- bias-dropout-add
class _BiasDropoutAdd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, residual, p, seed):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        r2 = residual.reshape(-1, x.shape[-1]).contiguous() if residual is not None else None
        out = torch.empty_like(x2)
        mask = torch.empty(x2.shape, dtype=torch.uint8, device=x.device) if p > 0 else None
        _lib.call("b2f_bias_dropout_add", ptr(x2), ptr(bias), ptr(r2), ptr(out), ptr(mask),
                  c_ll(x2.shape[0]), c_int(x2.shape[1]), c_float(p), c_ull(seed), stream_ptr())
        ctx.p, ctx.shape = p, x.shape
        ctx.has_bias, ctx.has_res = bias is not None, residual is not None
        if mask is not None:
            ctx.save_for_backward(mask)
        return out.reshape(x.shape)

    @staticmethod
    def backward(ctx, dout):
        d2 = dout.reshape(-1, dout.shape[-1]).contiguous()
        if ctx.p > 0:
            (mask,) = ctx.saved_tensors
            dx = torch.empty_like(d2)
            _lib.call("b2f_dropout_bwd", ptr(d2), ptr(mask), ptr(dx), c_ll(d2.numel()),
                      c_float(ctx.p), stream_ptr())
        else:
            dx = d2
        return (dx.reshape(ctx.shape), dx.sum(0) if ctx.has_bias else None,
                dout if ctx.has_res else None, None, None)


def bias_dropout_add(x, bias, residual, p: float = 0.0, seed: int = 0):
    """out = residual + dropout(x + bias, p)."""
    if not x.is_cuda:
        from . import ref
        return ref.bias_dropout_add(x, bias, residual, p, seed)
    return _BiasDropoutAdd.apply(x, bias, residual, float(p), int(seed))


# -------------------------------------------------------------------- RoPE
_rope_cache: dict = {}


def rope_tables(seq: int, dim: int, base: float, device) -> tuple[torch.Tensor, torch.Tensor]:
    key = (seq, dim, base, str(device))
    if key not in _rope_cache:
        inv = 1.0 / (base ** (torch.arange(0, dim, 2, dtype=torch.float32, device=device) / dim))
        ang = torch.outer(torch.arange(seq, dtype=torch.float32, device=device), inv)
        _rope_cache[key] = (ang.cos().contiguous(), ang.sin().contiguous())
    return _rope_cache[key]


def rope_inplace_raw(x: torch.Tensor, cos_t, sin_t, seq: int, backward: bool, pos_offset: int = 0,
                     pos_div: int = 1):
    """x: [tokens, heads, d] view with contiguous last two dims; rotated in place.
    Token t sits at position (t // pos_div) % seq + pos_offset."""
    tokens, nh, d = x.shape
    assert x.stride(2) == 1 and x.stride(1) == d
    _lib.call("b2f_rope", ptr(x), ptr(cos_t), ptr(sin_t), c_ll(tokens), c_int(seq), c_int(nh),
              c_int(d), c_ll(x.stride(0)), c_int(pos_offset), c_int(pos_div), c_int(int(backward)),
              stream_ptr())


class _Rope(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cos_t, sin_t, seq, pos_offset):
        ctx.mark_dirty(x)
        ctx.tables, ctx.seq, ctx.pos_offset = (cos_t, sin_t), seq, pos_offset
        rope_inplace_raw(x, cos_t, sin_t, seq, False, pos_offset)
        return x

    @staticmethod
    def backward(ctx, dx):
        dx = dx.contiguous() if not (dx.stride(2) == 1 and dx.stride(1) == dx.shape[2]) else dx
        rope_inplace_raw(dx, *ctx.tables, ctx.seq, True, ctx.pos_offset)
        return dx, None, None, None, None


def rope_(x, cos_t, sin_t, seq: int, pos_offset: int = 0):
    """In-place rotate-half rotary embedding of x[tokens, heads, d]."""
    return _Rope.apply(x, cos_t, sin_t, seq, pos_offset)


# ---------------------------------------------------------- cross-entropy
class _CrossEntropy(torch.autograd.Function):
    """Mean token loss; the logit gradient is produced by the forward kernel in place."""

    @staticmethod
    def forward(ctx, logits, targets, ignore_index):
        _lib.require_cuda(logits)
        l2 = logits.reshape(-1, logits.shape[-1])
        assert l2.stride(1) == 1
        t = targets.reshape(-1).to(torch.int64).contiguous()
        rows, V = l2.shape
        loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        valid = (t != ignore_index).sum().clamp(min=1).to(torch.float32)
        need_grad = logits.requires_grad
        # grad is written over the logits themselves (they are dead after the loss)
        _lib.call("b2f_cross_entropy", ptr(l2), ptr(t), ptr(loss), ptr(l2 if need_grad else None),
                  c_ll(rows), c_int(V), c_ll(l2.stride(0)), c_float(1.0), c_ll(ignore_index),
                  stream_ptr())
        ctx.save_for_backward(l2, valid)
        ctx.shape = logits.shape
        return loss.sum() / valid

    @staticmethod
    def backward(ctx, dloss):
        l2, valid = ctx.saved_tensors
        scale = (dloss / valid).to(torch.float32).reshape(1)
        _lib.call("b2f_scale", ptr(l2), c_ll(l2.numel()), c_int(0), ptr(scale), c_float(1.0),
                  stream_ptr())
        return l2.reshape(ctx.shape), None, None


def cross_entropy(logits, targets, ignore_index: int = -100):
    """Fused softmax-CE (mean over non-ignored tokens). Consumes `logits` (overwritten by grad)."""
    if not logits.is_cuda:
        from . import ref
        return ref.cross_entropy(logits, targets, ignore_index)
    return _CrossEntropy.apply(logits, targets, ignore_index)


def scaled_masked_softmax(x, mask=None, scale: float = 1.0, causal: bool = False):
    """x: [b, heads, sq, sk] bf16; mask: [b, sq, sk] uint8 (1 = masked out)."""
    b, h, sq, sk = x.shape
    x = x.contiguous()
    y = torch.empty_like(x)
    _lib.call("b2f_scaled_masked_softmax", ptr(x), ptr(mask), ptr(y), c_ll(b), c_int(h), c_int(sq),
              c_int(sk), c_float(scale), c_int(int(causal)), stream_ptr())
    return y


def cast_f32_bf16(x: torch.Tensor, out: torch.Tensor | None = None, scale: float = 1.0):
    if out is None:
        out = torch.empty(x.shape, dtype=_lib.HALF, device=x.device)
    _lib.call("b2f_cast_f32_bf16", ptr(x), ptr(out), c_ll(x.numel()), c_float(scale), stream_ptr())
    return out
