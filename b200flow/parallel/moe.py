"""Mixture-of-experts layer with expert parallelism.

What `deepspeed.moe.layer.MoE` gives the pinned Megatron-DeepSpeed (top-1 gate with capacity
factor and load-balancing loss -> einsum dispatch -> `all_to_all_single` -> local expert FFNs
-> `all_to_all_single` -> einsum combine; SURVEY.md §2F "EP", §2H C13-C15), re-designed:

* the router (K5a, csrc/kernels/moe.cu) emits a permutation, not dense one-hot operands;
* dispatch and combine are row movers; under expert parallelism they have two
  implementations: the *baseline* (`all_to_all_single` on [E, C, H] buffers, also the gloo
  test path) and the *fused* peer-memory path, where the dispatch kernel stores each token
  straight into the owning GPU's expert input buffer and the combine kernel pulls expert
  outputs back over NVLink, gate-scaled, into token order — the two all-to-alls disappear into
  the permutation kernels;
* the expert FFNs are one batched tcgen05 GEMM pair over the local experts.

Expert parameters are tagged `is_expert` so the engine keeps them in their own flat buffer
reduced over the expert-data-parallel group only.
"""
from __future__ import annotations

import ctypes
import math

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from ..ops import _lib
from ..ops import gemm as G
from ..ops._lib import c_int, ptr, stream_ptr


class EPContext:
    def __init__(self, group=None, world: int = 1, rank: int = 0, comm=None):
        self.group, self.world, self.rank, self.comm = group, world, rank, comm
        self._bufs: dict = {}
        self.calls = 0

    def symmetric(self, key, nbytes):
        k = (key, (nbytes + (1 << 20) - 1) >> 20)
        if k not in self._bufs:
            self._bufs[k] = [self.comm.alloc(k[1] << 20) for _ in range(2)]
        return self._bufs[k]

    def barrier(self):
        c = self.comm
        _lib.call("b2f_peer_barrier", c.flag_region(3), c_int(c.rank), c_int(c.world),
                  ctypes.c_uint(c.next_epoch()), stream_ptr())


LOCAL = EPContext()


# ------------------------------------------------------------------- router
def route_top1(logits: torch.Tensor, capacity: int):
    """logits fp32 [T, E] -> dict(expert, slot, gate, probs, counts, me, ce)."""
    T, E = logits.shape
    dev = logits.device
    if logits.is_cuda:
        out = dict(expert=torch.empty(T, dtype=torch.int32, device=dev), slot=torch.empty(T, dtype=torch.int32, device=dev),
                   gate=torch.empty(T, dtype=torch.float32, device=dev), probs=torch.empty(T, E, dtype=torch.float32, device=dev),
                   counts=torch.empty(E, dtype=torch.int32, device=dev), me=torch.empty(E, dtype=torch.float32, device=dev),
                   ce=torch.empty(E, dtype=torch.float32, device=dev))
        scratch = torch.empty(((T + 511) // 512) * E * 2, dtype=torch.int32, device=dev)   # per-block partials
        _lib.call("b2f_moe_router_top1", ptr(logits.contiguous()), c_int(T), c_int(E), c_int(capacity),
                  ptr(out["expert"]), ptr(out["slot"]), ptr(out["gate"]), ptr(out["probs"]), ptr(out["counts"]),
                  ptr(out["me"]), ptr(out["ce"]), ptr(scratch), stream_ptr())
        return out
    probs = logits.float().softmax(-1)
    gate, best = probs.max(-1)
    onehot = F.one_hot(best, E)
    slot = (onehot.cumsum(0) - 1).gather(1, best[:, None]).squeeze(1)
    keep = slot < capacity
    return dict(expert=torch.where(keep, best, -1).int(), slot=torch.where(keep, slot, -1).int(),
                gate=torch.where(keep, gate, 0.0), probs=probs, counts=onehot.sum(0).clamp(max=capacity).int(),
                me=probs.sum(0), ce=onehot.sum(0).float())


def route_top2(logits: torch.Tensor, capacity: int):
    """DeepSpeed `top2gating` without noise: every token goes to its two most probable experts.

    Second choices queue up behind *all* first choices of an expert (their slot starts at that
    expert's first-choice count), then the capacity cut is applied to both.  Built from two
    launches of the top-1 router (K5a) — the second on logits with the first choice masked out —
    plus a few [T]-sized index ops, so CUDA and CPU share the routing order.
    Returns dict(expert [2, T], slot [2, T] (-1 = dropped), probs [T, E], first_counts [E])."""
    T, E = logits.shape
    assert E >= 2, "top-2 routing needs at least two experts"
    r1 = route_top1(logits, T)                      # capacity T: nothing dropped, raw slots
    e1 = r1["expert"].long()
    masked = logits.masked_fill(F.one_hot(e1, E).bool(), torch.finfo(torch.float32).min)
    r2 = route_top1(masked, T)
    e2 = r2["expert"].long()
    first = r1["counts"].long()                     # first-choice tokens per expert (uncut)
    s1, s2 = r1["slot"].long(), r2["slot"].long() + first[e2]
    k1, k2 = s1 < capacity, s2 < capacity
    neg = torch.full_like(s1, -1)
    expert = torch.stack([torch.where(k1, e1, neg), torch.where(k2, e2, neg)]).int()
    slot = torch.stack([torch.where(k1, s1, neg), torch.where(k2, s2, neg)]).int()
    return dict(expert=expert, slot=slot, probs=r1["probs"], first_counts=first)


def gate_top2(logits: torch.Tensor, capacity: int):
    """(gates [2, T], aux, expert [2, T], slot [2, T]).  Routing is constant; the gate values
    g_k / (g_1 + g_2) (dropped choices count as 0) and the load-balancing loss
    E * sum_e mean_t(p_te) * frac_first_e are plain differentiable torch expressions of the
    softmax — [T, E] is tiny next to the expert GEMMs."""
    T, E = logits.shape
    with torch.no_grad():
        r = route_top2(logits.detach().float().contiguous(), capacity)
    probs = logits.float().softmax(-1)
    kept = r["expert"] >= 0
    g = probs.gather(1, r["expert"].clamp(min=0).long().t()).t() * kept       # [2, T]
    gates = g / g.sum(0, keepdim=True).clamp(min=torch.finfo(torch.float32).eps)
    aux = (probs.mean(0) * (r["first_counts"].float() / T)).sum() * E
    return gates, aux, r["expert"], r["slot"]


class _Gate(torch.autograd.Function):
    """Routing decisions are constants; gradients reach the logits through the chosen gate
    probability and the auxiliary loss  l_aux = E * sum_e mean_t(p_te) * frac_e."""

    @staticmethod
    def forward(ctx, logits, capacity):
        r = route_top1(logits.float(), capacity)
        T, E = logits.shape
        aux = (r["me"] / T * r["ce"] / T).sum() * E
        ctx.save_for_backward(r["probs"], r["expert"], r["ce"])
        ctx.mark_non_differentiable(r["expert"], r["slot"], r["counts"])
        return r["gate"], aux, r["expert"], r["slot"], r["counts"]

    @staticmethod
    def backward(ctx, dgate, daux, *_):
        probs, expert, ce = ctx.saved_tensors
        T, E = probs.shape
        # d l / d p[t, e]: chosen-expert gate prob + aux term E * ce_e / T^2
        dp = (daux * E / (T * T)) * ce[None, :].expand(T, E).clone()
        kept = expert >= 0
        idx = expert.clamp(min=0).long()
        dp[torch.arange(T, device=probs.device)[kept], idx[kept]] += dgate[kept]
        dlogits = probs * (dp - (dp * probs).sum(-1, keepdim=True))
        return dlogits, None


# ------------------------------------------------------- dispatch / combine
def _rows_index(expert, slot, capacity):
    return (expert.long() * capacity + slot.long()).clamp(min=0), expert >= 0


def _dispatch_local(x, expert, slot, E, capacity):
    T, H = x.shape
    buf = torch.zeros((E * capacity, H), dtype=x.dtype, device=x.device)
    if x.is_cuda:
        _lib.call("b2f_moe_dispatch", ptr(x), ptr(expert), ptr(slot), c_int(T), c_int(H), c_int(capacity),
                  ptr(buf), ptr(None), c_int(1), c_int(E), c_int(0), stream_ptr())
    else:
        rows, kept = _rows_index(expert, slot, capacity)
        buf[rows[kept]] = x[kept]
    return buf


def _combine_local(buf, expert, slot, gate, capacity):
    T = expert.shape[0]
    H = buf.shape[1]
    y = torch.empty((T, H), dtype=buf.dtype, device=buf.device)
    if buf.is_cuda:
        _lib.call("b2f_moe_combine", ptr(buf), ptr(None), c_int(1), c_int(1), c_int(0), ptr(expert), ptr(slot),
                  ptr(gate), c_int(T), c_int(H), c_int(capacity), ptr(y), stream_ptr())
    else:
        rows, kept = _rows_index(expert, slot, capacity)
        y = torch.where(kept[:, None], buf[rows] * gate[:, None].to(buf.dtype), torch.zeros((), dtype=buf.dtype))
    return y


def _a2a(buf, ep: EPContext):
    if ep.world == 1:
        return buf
    out = torch.empty_like(buf)
    if buf.is_cuda:
        dist.all_to_all_single(out, buf.contiguous(), group=ep.group)
    else:   # gloo has no all-to-all: gather every rank's buffer and keep the chunk addressed to us
        everyone = [torch.empty_like(buf) for _ in range(ep.world)]
        dist.all_gather(everyone, buf.contiguous(), group=ep.group)
        n = buf.shape[0] // ep.world
        for r in range(ep.world):
            out[r * n:(r + 1) * n] = everyone[r][ep.rank * n:(ep.rank + 1) * n]
    return out


class _Dispatch(torch.autograd.Function):
    """x [T, H] -> expert inputs [E_local, ep * C, H] on the owners (fused or baseline)."""

    @staticmethod
    def forward(ctx, x, expert, slot, E, capacity, ep: EPContext, fused):
        ctx.save_for_backward(expert, slot)
        ctx.cfg = (E, capacity, ep, fused, x.shape[0])
        return _dispatch_fwd(x.contiguous(), expert, slot, E, capacity, ep, fused)

    @staticmethod
    def backward(ctx, dbuf):
        expert, slot = ctx.saved_tensors
        E, capacity, ep, fused, T = ctx.cfg
        ones = torch.ones(T, dtype=torch.float32, device=dbuf.device)
        dx = _combine_fwd(dbuf.contiguous(), expert, slot, ones, E, capacity, ep, fused)
        return dx, None, None, None, None, None, None


class _Combine(torch.autograd.Function):
    """expert outputs [E_local, ep * C, H] -> y[t] = gate[t] * out_row(t)."""

    @staticmethod
    def forward(ctx, out, gate, expert, slot, E, capacity, ep: EPContext, fused):
        out = out.contiguous()
        y, rows_local = _combine_fwd(out, expert, slot, gate, E, capacity, ep, fused, return_rows=True)
        ctx.save_for_backward(expert, slot, gate, rows_local)
        ctx.cfg = (E, capacity, ep, fused)
        return y

    @staticmethod
    def backward(ctx, dy):
        expert, slot, gate, rows_local = ctx.saved_tensors
        E, capacity, ep, fused = ctx.cfg
        dy = dy.contiguous()
        T, H = dy.shape
        if dy.is_cuda:
            dgate = torch.empty(T, dtype=torch.float32, device=dy.device)
            _lib.call("b2f_moe_gate_grad", ptr(rows_local), ptr(expert), ptr(slot), ptr(dy), c_int(T), c_int(H),
                      c_int(capacity), ptr(dgate), stream_ptr())
        else:
            rows, kept = _rows_index(expert, slot, capacity)
            dgate = torch.where(kept, (rows_local[rows].float() * dy.float()).sum(-1), torch.zeros(()))
        dout = _dispatch_fwd((dy * gate[:, None].to(dy.dtype)).contiguous(), expert, slot, E, capacity, ep, fused)
        return dout, dgate, None, None, None, None, None, None


def _dispatch_fwd(x, expert, slot, E, capacity, ep: EPContext, fused):
    T, H = x.shape
    El = E // ep.world
    if ep.world == 1:
        return _dispatch_local(x, expert, slot, E, capacity).view(E, capacity, H)
    if fused and x.is_cuda:
        ep.calls += 1
        recv = ep.symmetric("in", ep.world * El * capacity * H * 2)[ep.calls & 1]
        view = recv.view(_lib.HALF, ep.world * El * capacity * H)
        view.zero_()
        ep.barrier()                      # everyone's buffer is cleared before anyone stores into it
        _lib.call("b2f_moe_dispatch", ptr(x), ptr(expert), ptr(slot), c_int(T), c_int(H), c_int(capacity),
                  ptr(None), recv.ptr_array, c_int(ep.world), c_int(El), c_int(ep.rank), stream_ptr())
        ep.barrier()                      # every peer's rows have landed
        # copy out of the (reused) symmetric buffer: the expert GEMMs save their input for backward
        got = view.view(ep.world, El, capacity, H).clone()
    else:
        buf = _dispatch_local(x, expert, slot, E, capacity)            # [E*C, H] = [ep, El, C, H]
        got = _a2a(buf, ep).view(ep.world, El, capacity, H)
    return got.permute(1, 0, 2, 3).reshape(El, ep.world * capacity, H)


def _combine_fwd(out, expert, slot, gate, E, capacity, ep: EPContext, fused, return_rows=False):
    El, _, H = out.shape
    T = expert.shape[0]
    if ep.world == 1:
        rows = out.reshape(E * capacity, H)
        y = _combine_local(rows, expert, slot, gate, capacity)
        return (y, rows) if return_rows else y
    back = out.view(El, ep.world, capacity, H).permute(1, 0, 2, 3).contiguous()    # [ep(src), El, C, H]
    if fused and out.is_cuda:
        ep.calls += 1
        sym = ep.symmetric("out", back.numel() * 2)[ep.calls & 1]
        sym.view(_lib.HALF, back.numel()).copy_(back.reshape(-1))
        ep.barrier()                      # all owners published their expert outputs
        y = torch.empty((T, H), dtype=out.dtype, device=out.device)
        _lib.call("b2f_moe_combine", ptr(None), sym.ptr_array, c_int(ep.world), c_int(El), c_int(ep.rank),
                  ptr(expert), ptr(slot), ptr(gate), c_int(T), c_int(H), c_int(capacity), ptr(y), stream_ptr())
        ep.barrier()                      # nobody overwrites its buffer while peers still read it
        if not return_rows:
            return y
        # rows of this rank's own tokens (needed for d gate): pull with unit gate
        ones = torch.ones(T, dtype=torch.float32, device=out.device)
        rows = torch.zeros((E * capacity, H), dtype=out.dtype, device=out.device)
        mine = torch.empty((T, H), dtype=out.dtype, device=out.device)
        _lib.call("b2f_moe_combine", ptr(None), sym.ptr_array, c_int(ep.world), c_int(El), c_int(ep.rank),
                  ptr(expert), ptr(slot), ptr(ones), c_int(T), c_int(H), c_int(capacity), ptr(mine), stream_ptr())
        ep.barrier()
        r, kept = _rows_index(expert, slot, capacity)
        rows[r[kept]] = mine[kept]
        return y, rows
    rows = _a2a(back.view(E * capacity, H), ep)                                      # my tokens' outputs
    y = _combine_local(rows, expert, slot, gate, capacity)
    return (y, rows) if return_rows else y


# --------------------------------------------------------------------- layer
class MoELayer(nn.Module):
    def __init__(self, hidden: int, ffn: int, num_experts: int, ep: EPContext | None = None,
                 capacity_factor: float = 1.0, min_capacity: int = 4, act: str = "gelu",
                 aux_loss_coeff: float = 0.01, init_std: float = 0.02, dtype=torch.bfloat16, device=None,
                 fused: bool = True, fp8: bool = False, top_k: int = 1, eval_capacity_factor: float | None = None):
        super().__init__()
        assert top_k in (1, 2), "DeepSpeed's TopKGate supports k = 1 or 2"
        self.fp8, self.top_k = fp8, top_k
        self.ep = ep or LOCAL
        assert num_experts % self.ep.world == 0
        self.E, self.El = num_experts, num_experts // self.ep.world
        self.hidden, self.ffn, self.cf, self.min_cap = hidden, ffn, capacity_factor, min_capacity
        self.eval_cf = capacity_factor if eval_capacity_factor is None else eval_capacity_factor   # --moe-eval-capacity-factor
        self.aux_coeff, self.fused = aux_loss_coeff, fused
        kw = dict(dtype=dtype, device=device)
        e_pad = (num_experts + 7) // 8 * 8
        self.wg = nn.Parameter(torch.empty(e_pad, hidden, **kw).normal_(0, init_std))
        self.w1 = nn.Parameter(torch.empty(self.El, ffn, hidden, **kw).normal_(0, init_std))
        self.b1 = nn.Parameter(torch.zeros(self.El, ffn, **kw))
        self.w2 = nn.Parameter(torch.empty(self.El, hidden, ffn, **kw).normal_(0, init_std))
        self.b2 = nn.Parameter(torch.zeros(self.El, hidden, **kw))
        for p in (self.w1, self.b1, self.w2, self.b2):
            p.is_expert = True
        self.last_aux = None
        self.last_counts = None

    def capacity(self, tokens: int) -> int:
        cf = self.cf if self.training else self.eval_cf
        c = max(self.min_cap, math.ceil(tokens / self.E * cf * self.top_k))
        return (c + 7) // 8 * 8

    def _use_k5(self, x, C) -> bool:
        import os

        from . import fused_moe as FM

        if not (self.fp8 and x.is_cuda and self.fused and os.environ.get("B2F_FUSED_MOE", "1") == "1"):
            return False
        if self.ep.world > 1 and self.ep.comm is None:
            return False
        return FM.supported(self.hidden, self.ffn, C, x.shape[0])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        T, H = x.shape
        C = self.capacity(T)
        if x.is_cuda:
            logits = _GateLogits.apply(x, self.wg, self.E)
        else:
            logits = G.linear(x, self.wg)[:, :self.E]
        if self.top_k == 2:
            # both choices ride through the top-1 row movers as a token list of length 2T
            gates, aux, expert2, slot2 = gate_top2(logits, C)
            gate, expert, slot = gates.reshape(-1), expert2.reshape(-1).contiguous(), slot2.reshape(-1).contiguous()
            counts = torch.bincount(expert.clamp(min=0).long(), weights=(expert >= 0).float(),
                                    minlength=self.E).int()
            x = torch.cat([x, x], 0)
        else:
            gate, aux, expert, slot, counts = _Gate.apply(logits.float(), C)
        self.last_aux, self.last_counts = aux, counts
        if self._use_k5(x, C):
            # K5: dispatch + expert GEMM 1 and expert GEMM 2 + combine as two fused block-scaled-fp8 kernels
            from . import fused_moe as FM

            if getattr(self.ep, "fused_moe", None) is None:
                self.ep.fused_moe = FM.FusedMoE(self.ep)
            y = FM.fused_experts(x, gate, self.w1, self.b1, self.w2, self.b2, expert, slot, counts, self.E, C,
                                 self.ep.fused_moe)
            return y[:T] + y[T:] if self.top_k == 2 else y
        xin = _Dispatch.apply(x, expert, slot, self.E, C, self.ep, self.fused)       # [El, ep*C, H]
        outs = []
        if self.fp8 and x.is_cuda:
            from ..ops.gemm_fp8 import linear_fp8 as lin, linear_gelu_fp8 as lin_gelu
        else:
            lin, lin_gelu = G.linear, G.linear_gelu
        for e in range(self.El):
            h1 = lin_gelu(xin[e], self.w1[e], self.b1[e])
            outs.append(lin(h1, self.w2[e], self.b2[e]))
        out = torch.stack(outs, 0)
        y = _Combine.apply(out, gate, expert, slot, self.E, C, self.ep, self.fused)
        if self.top_k == 2:
            y = y[:T] + y[T:]
        return y


class _GateLogits(torch.autograd.Function):
    """fp32 gate logits on the tcgen05 GEMM (the gate weight is padded to 8 rows)."""

    @staticmethod
    def forward(ctx, x, wg, E):
        ctx.save_for_backward(x, wg)
        ctx.E = E
        return G.gemm(x.contiguous(), wg, out_dtype=torch.float32)[:, :E]

    @staticmethod
    def backward(ctx, dl):
        x, wg = ctx.saved_tensors
        d = torch.zeros((x.shape[0], wg.shape[0]), dtype=_lib.HALF, device=x.device)
        d[:, :ctx.E] = dl.to(_lib.HALF)
        dx = G.gemm(d, wg, b_mn_major=True)
        dw = G._wgrad(x, d, wg)
        return dx, dw, None
