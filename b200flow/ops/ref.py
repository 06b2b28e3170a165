"""Plain-PyTorch fp32-accurate oracles of every native op.

Two uses: (1) numerics tests compare the sm_100a kernels against these;
(2) CPU tensors are routed here so that the launcher / parallel-runtime plumbing
(gloo, BASELINE.json config 1) runs on machines without a GPU.  CUDA tensors never
take this path — on a GPU box a missing native library is an error, not a fallback.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def linear(x, weight, bias=None, residual=None):
    y = F.linear(x, weight, bias)
    return y + residual if residual is not None else y


def linear_gelu(x, weight, bias):
    return F.gelu(F.linear(x, weight, bias), approximate="tanh")


def norm(x, gamma, beta=None, residual=None, eps=1e-5, rms=False):
    h = x + residual if residual is not None else x
    hf = h.float()
    if rms:
        y = hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps) * gamma.float()
    else:
        y = F.layer_norm(hf, (h.shape[-1],), gamma.float(), beta.float() if beta is not None else None, eps)
    y = y.to(x.dtype)
    return (y, h) if residual is not None else y


def swiglu(gu):
    f = gu.shape[-1] // 2
    return F.silu(gu[..., :f]) * gu[..., f:]


def bias_gelu(x, bias=None):
    return F.gelu(x + bias if bias is not None else x, approximate="tanh")


def bias_dropout_add(x, bias, residual, p=0.0, seed=0, training=True):
    y = x + bias if bias is not None else x
    if p > 0 and training:      # the mask is a function of the seed alone, like the kernel's counter-based one
        g = torch.Generator().manual_seed(int(seed) & 0x7FFFFFFFFFFFFFFF)
        keep = (torch.rand(y.shape, generator=g) >= p).to(y.dtype)
        y = y * keep / (1.0 - p)
    return y + residual if residual is not None else y


def rope(x, cos_t, sin_t, positions):
    """x: [tokens, heads, d]; positions: [tokens] int64."""
    half = x.shape[-1] // 2
    c = cos_t[positions][:, None, :].to(torch.float32)
    s = sin_t[positions][:, None, :].to(torch.float32)
    x1, x2 = x[..., :half].float(), x[..., half:].float()
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], -1).to(x.dtype)


def attention(q, k, v, causal=True, scale=None, kv_lens=None, dropout_p=0.0, seed=0):
    """q: [b, s, h, d]; k, v: [b, s, hkv, d] -> [b, s, h, d]."""
    B, S, H, D = q.shape
    HKV = k.shape[2]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    qf, kf, vf = q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float()
    if HKV != H:
        kf = kf.repeat_interleave(H // HKV, 1)
        vf = vf.repeat_interleave(H // HKV, 1)
    s = qf @ kf.transpose(-1, -2) * scale
    mask = torch.zeros(B, 1, S, S, dtype=torch.bool, device=q.device)
    if causal:
        mask = mask | torch.triu(torch.ones(S, S, dtype=torch.bool, device=q.device), 1)
    if kv_lens is not None:
        mask = mask | (torch.arange(S, device=q.device)[None, :] >= kv_lens[:, None])[:, None, None, :]
    s = s.masked_fill(mask, float("-inf"))
    pr = torch.softmax(s, -1)
    if dropout_p > 0:      # the kernels' counter-based mask, restated (ops/attention.py)
        from .attention import dropout_keep_mask

        keep, rinv = dropout_keep_mask(B, H, S, dropout_p, seed, q.device)
        pr = pr * keep * rinv
    return (pr @ vf).transpose(1, 2).to(q.dtype)


def cross_entropy(logits, targets, ignore_index=-100):
    return F.cross_entropy(logits.float().reshape(-1, logits.shape[-1]), targets.reshape(-1),
                           ignore_index=ignore_index)


def adamw_step(p, m, v, g, *, lr, beta1, beta2, eps, weight_decay, step):
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    mhat = m / (1 - beta1 ** step)
    vhat = v / (1 - beta2 ** step)
    p.add_(mhat / (vhat.sqrt() + eps) + weight_decay * p, alpha=-lr)
