"""fp8 (e4m3) linear layers on the tcgen05 `kind::f8f6f4` GEMM (csrc/kernels/gemm.cu).

Recipe: activations are quantised per token (row), weights per output
channel, both to e4m3 with scale = amax / 448; the product accumulates in fp32 in TMEM and the
epilogue multiplies by row_scale[m] * col_scale[n], which is exact because both scales are
constant along the contraction.  Forward and dgrad run in fp8 (dgrad re-quantises dy per row
and the transposed weight per input channel); wgrad stays bf16 with fp32 accumulation, since
per-token scales do not factor out of a contraction over tokens.

Used by the MoE experts for BASELINE.json config 5 ("GPT-MoE 1.3B ... fp8").
"""
from __future__ import annotations

import torch

from . import _lib
from . import gemm as G
from ._lib import c_int, c_ll, ptr, stream_ptr


def quantize_rows(x: torch.Tensor):
    """x bf16 [R, K] -> (q uint8 [R, K] holding e4m3 bits, scale fp32 [R])."""
    assert x.dim() == 2 and x.dtype == _lib.HALF and x.stride(1) == 1
    R, K = x.shape
    q = torch.empty((R, K), dtype=torch.uint8, device=x.device)
    s = torch.empty(R, dtype=torch.float32, device=x.device)
    _lib.call("b2f_quantize_rows_e4m3", ptr(x), ptr(q), ptr(s), c_ll(R), c_int(K), c_ll(x.stride(0)),
              stream_ptr())
    return q, s


def gemm_fp8(aq, a_scale, bq, b_scale, bias=None, residual=None, aux=None, gelu=False,
             out_dtype=None):
    """(aq [M,K] x bq [N,K]^T) * a_scale[m] * b_scale[n] (+bias) (gelu) (+residual)."""
    M, K = aq.shape
    N = bq.shape[0]
    out_dtype = out_dtype or _lib.HALF
    out = torch.empty((M, N), dtype=out_dtype, device=aq.device)
    flags = (G.EPI_GELU if gelu else 0) | (G.EPI_OUT_F32 if out_dtype == torch.float32 else 0)
    _lib.call("b2f_gemm_fp8", ptr(aq), ptr(bq), ptr(out), c_int(M), c_int(N), c_int(K), c_ll(aq.stride(0)),
              c_ll(bq.stride(0)), c_ll(out.stride(0)), ptr(a_scale), ptr(b_scale), ptr(bias), ptr(residual),
              c_ll(residual.stride(0) if residual is not None else 0), ptr(aux),
              c_ll(aux.stride(0) if aux is not None else 0), c_int(flags), stream_ptr())
    return out


class _LinearFP8(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, gelu):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        xq, sx = quantize_rows(x2)
        wq, sw = quantize_rows(weight)
        pre = torch.empty((x2.shape[0], weight.shape[0]), dtype=_lib.HALF, device=x.device) if gelu else None
        y = gemm_fp8(xq, sx, wq, sw, bias=bias, aux=pre, gelu=gelu)
        ctx.save_for_backward(x2, weight, pre if gelu else x2.new_empty(0))
        ctx.gelu, ctx.has_bias, ctx.shape = gelu, bias is not None, x.shape
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        from .elementwise import gelu_bwd_raw

        x2, weight, pre = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        if ctx.gelu:
            dy2 = gelu_bwd_raw(pre, None, dy2)
        dq, sd = quantize_rows(dy2)
        wtq, swt = quantize_rows(weight.t().contiguous())
        dx = gemm_fp8(dq, sd, wtq, swt).reshape(ctx.shape)
        dw = G._wgrad(x2, dy2, weight)
        return dx, dw, (dy2.sum(0) if ctx.has_bias else None), None


def linear_fp8(x, weight, bias=None):
    if not x.is_cuda:
        from . import ref
        return ref.linear(x, weight, bias)
    return _LinearFP8.apply(x, weight, bias, False)


def linear_gelu_fp8(x, weight, bias=None):
    if not x.is_cuda:
        from . import ref
        return ref.linear_gelu(x, weight, bias)
    return _LinearFP8.apply(x, weight, bias, True)
