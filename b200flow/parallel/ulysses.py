"""Ulysses-style sequence parallelism for attention (long context).

Outside attention every rank of the sequence-parallel group holds a contiguous slice of the
sequence (`[b, s/P, h, d]`).  Attention needs whole sequences, so around it the group swaps the
partition from *sequence* to *heads* with one all-to-all, runs the ordinary flash attention kernel on
`[b, s, h/P, d]`, and swaps back.  This is what DeepSpeed-Ulysses (`DistributedAttention`) does in the
Megatron-DeepSpeed the reference pins (its `--ds-sequence-parallel-size`; the reference's example
values files leave it at 1 and rely on Megatron sequence parallelism, which `tensor_parallel.py`
covers).  Backward of each swap is the opposite swap.

The all-to-all is NCCL `all_to_all_single` on GPUs and an all-gather based emulation on gloo (CPU
tests).  K and V with fewer heads than the group size (GQA) are replicated up to the group size first.
"""
from __future__ import annotations

import math

import torch
import torch.distributed as dist

from ..ops import attention as A


def _all_to_all(chunks_in: torch.Tensor, group) -> torch.Tensor:
    """chunks_in: [P, ...] (chunk r goes to rank r) -> [P, ...] (chunk r came from rank r)."""
    world = dist.get_world_size(group)
    out = torch.empty_like(chunks_in)
    if chunks_in.is_cuda:
        dist.all_to_all_single(out, chunks_in.contiguous(), group=group)
        return out
    rank = dist.get_rank(group)
    everyone = [torch.empty_like(chunks_in) for _ in range(world)]
    dist.all_gather(everyone, chunks_in.contiguous(), group=group)
    for r in range(world):
        out[r] = everyone[r][rank]
    return out


def _seq_to_heads(x: torch.Tensor, group) -> torch.Tensor:
    """[b, s/P, h, d] -> [b, s, h/P, d]."""
    P = dist.get_world_size(group)
    b, sl, h, d = x.shape
    assert h % P == 0, "heads must divide by the sequence-parallel group size"
    send = x.reshape(b, sl, P, h // P, d).permute(2, 0, 1, 3, 4).contiguous()      # [P, b, s/P, h/P, d]
    recv = _all_to_all(send, group)                                                # chunk r = rank r's slice
    return recv.permute(1, 0, 2, 3, 4).reshape(b, P * sl, h // P, d)


def _heads_to_seq(x: torch.Tensor, group) -> torch.Tensor:
    """[b, s, h/P, d] -> [b, s/P, h, d]."""
    P = dist.get_world_size(group)
    b, s, hl, d = x.shape
    sl = s // P
    send = x.reshape(b, P, sl, hl, d).permute(1, 0, 2, 3, 4).contiguous()          # [P, b, s/P, h/P, d]
    recv = _all_to_all(send, group)                                                # chunk r = head group r
    return recv.permute(1, 2, 0, 3, 4).reshape(b, sl, P * hl, d)


class _SeqToHeads(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return _seq_to_heads(x, group)

    @staticmethod
    def backward(ctx, g):
        return _heads_to_seq(g.contiguous(), ctx.group), None


class _HeadsToSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return _heads_to_seq(x, group)

    @staticmethod
    def backward(ctx, g):
        return _seq_to_heads(g.contiguous(), ctx.group), None


def ulysses_attention(q, k, v, group=None, causal: bool = True, scale: float | None = None):
    """q: [b, s/P, h, d], k/v: [b, s/P, hkv, d] sequence shards -> output shard [b, s/P, h, d]."""
    if group is None or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return A.flash_attention(q, k, v, causal=causal, scale=scale)
    P = dist.get_world_size(group)
    if scale is None:
        scale = 1.0 / math.sqrt(q.shape[-1])
    if k.shape[2] % P != 0:      # GQA with fewer kv heads than ranks: replicate up to a multiple of P
        rep = P // math.gcd(k.shape[2], P)
        k = k.repeat_interleave(rep, dim=2)
        v = v.repeat_interleave(rep, dim=2)
    qh = _SeqToHeads.apply(q, group)
    kh = _SeqToHeads.apply(k, group)
    vh = _SeqToHeads.apply(v, group)
    oh = A.flash_attention(qh, kh, vh, causal=causal, scale=scale)
    return _HeadsToSeq.apply(oh.contiguous(), group)


class _AllToAll(torch.autograd.Function):
    """Chunk r of the leading dimension goes to rank r; the gradient takes the same route back."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return _all_to_all(x.contiguous(), group)

    @staticmethod
    def backward(ctx, g):
        return _all_to_all(g.contiguous(), ctx.group), None


def ulysses_qkv_attention(qkv, heads: int, kv_heads: int, causal: bool, scale: float, group, attention_fn,
                          rope_tables=None, kv_lens=None):
    """The engine's form of DeepSpeed-Ulysses (`--ds-sequence-parallel-size`).

    qkv: this rank's sequence slice of the fused projection, `[s/P, b, heads + 2 kv_heads, d]` (sequence-major,
    slice r = positions r s/P ...).  One all-to-all regroups it to the whole sequence of `heads/P + 2 kv_heads/P`
    heads — still the packed layout, so the ordinary fused RoPE + flash-attention node (`attention_fn`) runs on it
    unchanged, with positions counted over the full sequence — and a second one brings the output back to
    `[s/P * b, heads * d]`."""
    P = dist.get_world_size(group)
    sl, b, _, d = qkv.shape
    assert heads % P == 0 and kv_heads % P == 0, "query and key/value heads must divide by the sequence-parallel size"
    hl, kl = heads // P, kv_heads // P
    q, k, v = qkv[:, :, :heads], qkv[:, :, heads:heads + kv_heads], qkv[:, :, heads + kv_heads:]
    send = torch.cat([q.reshape(sl, b, P, hl, d), k.reshape(sl, b, P, kl, d), v.reshape(sl, b, P, kl, d)], dim=3)
    recv = _AllToAll.apply(send.permute(2, 0, 1, 3, 4), group)                 # [P (sequence slice), s/P, b, nh/P, d]
    full = recv.reshape(P * sl, b, hl + 2 * kl, d)
    o = attention_fn(full, hl, kl, causal, scale, rope_tables, kv_lens)          # [s * b, hl * d]
    back = _AllToAll.apply(o.reshape(P, sl, b, hl, d), group)                    # [P (head group), s/P, b, hl, d]
    return back.permute(1, 2, 0, 3, 4).reshape(sl * b, heads * d)
