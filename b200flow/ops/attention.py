"""K7 — tcgen05 flash attention (csrc/kernels/attention.cu) as an autograd function.

Tensors are ``[batch, seq, heads, head_dim]`` views whose last two dims are packed
(any token / batch stride), e.g. slices of a fused QKV projection output.
"""
from __future__ import annotations

import ctypes
import math

import torch

from . import _lib
from ._lib import c_float, c_int, c_ll, ptr, stream_ptr


def _strides(t: torch.Tensor):
    B, S, H, D = t.shape
    assert t.stride(3) == 1 and t.stride(2) == D, "heads must be packed: [b, s, h, d] with h*d contiguous"
    return t.stride(1), t.stride(0)


def dropout_keep_mask(B: int, H: int, S: int, p: float, seed: int, device="cpu") -> tuple[torch.Tensor, float]:
    """The attention kernels' dropout mask restated in PyTorch: (keep [B, H, S, S] bool, scale).  One 32-bit hash
    word per (batch * H + head, query, key // 4), one byte per key; an entry is kept iff its byte >= round(256 p),
    kept entries are scaled by 256 / (256 - round(256 p)).  Oracle for the tests and the CPU path."""
    M = 0xFFFFFFFF
    thr = int(p * 256.0 + 0.5)

    def mix(x):
        x = x ^ (x >> 16)
        x = (x * 0x7FEB352D) & M
        x = x ^ (x >> 15)
        x = (x * 0x846CA68B) & M
        return x ^ (x >> 16)

    bh = torch.arange(B * H, dtype=torch.int64, device=device)
    seed_bh = mix(((seed & M) + bh * 0x632BE5AB) & M)                              # [BH]
    q = torch.arange(S, dtype=torch.int64, device=device)
    seed_q = (seed_bh[:, None] + q[None, :] * 0x9E3779B1) & M                      # [BH, S]
    k = torch.arange(S, dtype=torch.int64, device=device)
    word = mix((seed_q[:, :, None] + (k >> 2)[None, None, :] * 0x85EBCA77) & M)    # [BH, S, S]
    byte = (word >> ((k & 3) * 8)[None, None, :]) & 0xFF
    return (byte >= thr).view(B, H, S, S), 256.0 / (256.0 - thr)


def attn_fwd_raw(q, k, v, causal: bool, scale: float, kv_lens=None, dropout_p: float = 0.0, seed: int = 0):
    B, S, H, D = q.shape
    HKV = k.shape[2]
    o = torch.empty((B, S, H, D), dtype=_lib.HALF, device=q.device)
    lse = torch.empty((B, H, S), dtype=torch.float32, device=q.device)
    (q_ts, q_bs), (k_ts, k_bs), (v_ts, v_bs) = _strides(q), _strides(k), _strides(v)
    args = [ptr(q), ptr(k), ptr(v), ptr(o), ptr(lse), c_int(B), c_int(S), c_int(H),
            c_int(HKV), c_int(D), c_ll(q_ts), c_ll(q_bs), c_ll(k_ts), c_ll(k_bs), c_ll(v_ts),
            c_ll(v_bs), c_ll(o.stride(1)), c_ll(o.stride(0)), c_float(scale), c_int(int(causal)), ptr(kv_lens)]
    if dropout_p > 0:
        _lib.call("b2f_attn_fwd_drop", *args, c_float(dropout_p), ctypes.c_uint(seed & 0xFFFFFFFF), stream_ptr())
    else:
        _lib.call("b2f_attn_fwd", *args, stream_ptr())
    _lib.debug_sync()
    return o, lse


def attn_bwd_raw(q, k, v, o, do, lse, causal: bool, scale: float, kv_lens=None, dq=None, dk=None,
                 dv=None, dropout_p: float = 0.0, seed: int = 0):
    B, S, H, D = q.shape
    HKV = k.shape[2]
    if dq is None:
        dq = torch.empty((B, S, H, D), dtype=_lib.HALF, device=q.device)
    if dk is None:
        dk = torch.empty((B, S, HKV, D), dtype=_lib.HALF, device=q.device)
    if dv is None:
        dv = torch.empty((B, S, HKV, D), dtype=_lib.HALF, device=q.device)
    s_pad = (S + 127) // 128 * 128
    # scratch for the pre-pass: [lse * log2e | rowsum(o * do)], query rows padded to whole tiles
    delta = torch.empty((2, B, H, s_pad), dtype=torch.float32, device=q.device)
    if not (do.stride(3) == 1 and do.stride(2) == D):
        do = do.contiguous()
    args = []
    for t in (q, k, v, o, do, dq, dk, dv):
        ts, bs = _strides(t)
        args += [c_ll(ts), c_ll(bs)]
    head = [ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(delta), ptr(dq), ptr(dk), ptr(dv), c_int(B),
            c_int(S), c_int(H), c_int(HKV), c_int(D), *args, c_float(scale), c_int(int(causal)), ptr(kv_lens)]
    if dropout_p > 0:
        _lib.call("b2f_attn_bwd_drop", *head, c_float(dropout_p), ctypes.c_uint(seed & 0xFFFFFFFF), stream_ptr())
    else:
        _lib.call("b2f_attn_bwd", *head, stream_ptr())
    _lib.debug_sync()
    return dq, dk, dv


class _FlashAttn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, causal, scale, kv_lens, dropout_p=0.0, seed=0):
        _lib.require_cuda(q, k, v)
        o, lse = attn_fwd_raw(q, k, v, causal, scale, kv_lens, dropout_p, seed)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.causal, ctx.scale, ctx.kv_lens, ctx.drop = causal, scale, kv_lens, (dropout_p, seed)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        dq, dk, dv = attn_bwd_raw(q, k, v, o, do, lse, ctx.causal, ctx.scale, ctx.kv_lens,
                                  dropout_p=ctx.drop[0], seed=ctx.drop[1])
        return dq, dk, dv, None, None, None, None, None


def flash_attention(q, k, v, causal: bool = True, scale: float | None = None, kv_lens=None,
                    dropout_p: float = 0.0, seed: int = 0):
    """softmax(scale * q kᵀ [+ causal / key-length mask]) v, O(seq) memory.  `dropout_p` > 0 drops softmax
    probabilities (Megatron's --attention-dropout) with a mask that is a function of (seed, batch, head, query, key)."""
    if scale is None:
        scale = 1.0 / math.sqrt(q.shape[-1])
    if not q.is_cuda:
        from . import ref
        return ref.attention(q, k, v, causal, scale, kv_lens, dropout_p, seed)
    return _FlashAttn.apply(q, k, v, causal, float(scale), kv_lens, float(dropout_p), int(seed))


class _FlashAttnPackedQKV(torch.autograd.Function):
    """qkv: [b, s, 3, h, d] (one GEMM output). The backward writes dq/dk/dv straight into one
    [b, s, 3, h, d] buffer so the QKV-projection dgrad/wgrad GEMMs consume it without a concat."""

    @staticmethod
    def forward(ctx, qkv, causal, scale, kv_lens):
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        o, lse = attn_fwd_raw(q, k, v, causal, scale, kv_lens)
        ctx.save_for_backward(qkv, o, lse)
        ctx.causal, ctx.scale, ctx.kv_lens = causal, scale, kv_lens
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse = ctx.saved_tensors
        dqkv = torch.empty_like(qkv)
        attn_bwd_raw(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], o, do, lse, ctx.causal, ctx.scale,
                     ctx.kv_lens, dq=dqkv[:, :, 0], dk=dqkv[:, :, 1], dv=dqkv[:, :, 2])
        return dqkv, None, None, None


def flash_attention_qkvpacked(qkv, causal: bool = True, scale: float | None = None, kv_lens=None):
    if scale is None:
        scale = 1.0 / math.sqrt(qkv.shape[-1])
    if not qkv.is_cuda:
        from . import ref
        return ref.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal, scale, kv_lens)
    return _FlashAttnPackedQKV.apply(qkv, causal, float(scale), kv_lens)
