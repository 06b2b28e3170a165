"""Megatron tensor parallelism with sequence parallelism.

`TPContext` is the object every linear of the model goes through:

* ``column_linear``  : Y = AG(X_shard) · W_colᵀ      (QKV, MLP up / gate)
* ``row_linear``     : Y_shard = RS(X · W_rowᵀ) + residual_shard   (attn-out, MLP down)
* ``scatter_sequence`` / ``lm_head_loss`` (vocab-parallel cross-entropy)

On one GPU these are plain tcgen05 GEMMs.  Across a TP group there are two
implementations behind the same interface: the *baseline* path (NCCL/gloo
collective followed by / preceded by the GEMM — what Megatron does, SURVEY.md
§2F/§2H C7-C10) and the *fused* path (`parallel/fused_tp.py`: K3 all-gather→GEMM,
K4 GEMM→reduce-scatter and K4′ GEMM→all-reduce kernels over NVLink peer memory).  The baseline path
also runs on CPU/gloo, which is how the sharding math is unit-tested.

With ``sequence_parallel=False`` the context reproduces W3's non-SP layout
(`examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml:39-55`):
identity-forward/all-reduce-backward before column linears and
all-reduce-forward after row linears; on GPUs both all-reduces are fused into the
producing GEMM (K4′) when the token count is a multiple of 128 x TP.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ..ops import elementwise as E
from ..ops import _lib
from ..ops import gemm as G


_FUSED_AR = os.environ.get("B2F_FUSED_AR", "1") == "1"   # B2F_FUSED_AR=0: NCCL all-reduce around the GEMM


# ---------------------------------------------------------------- primitives


def _mm(x, w, bias=None, residual=None):
    """x [M,K] · w[N,K]ᵀ (+bias) (+residual)."""
    if x.is_cuda:
        return G.gemm(x, w, bias=bias, residual=residual)
    y = F.linear(x, w, bias)
    return y + residual if residual is not None else y


def _mm_dgrad(dy, w):
    if dy.is_cuda:
        return G.gemm(dy, w, b_mn_major=True)
    return dy @ w


def _mm_wgrad(x, dy, w):
    """Returns dW, or None when accumulated into w.main_grad by the GEMM epilogue."""
    if dy.is_cuda or (G._DEFERRED is not None and getattr(w, "main_grad", None) is not None):
        return G._wgrad(x, dy, w)
    return dy.t() @ x


def _all_gather_rows(x, group, world):
    out = torch.empty((x.shape[0] * world, *x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x.contiguous(), group=group)
    return out


def _reduce_scatter_rows(x, group, world):
    out = torch.empty((x.shape[0] // world, *x.shape[1:]), dtype=x.dtype, device=x.device)
    if x.is_cuda:
        dist.reduce_scatter_tensor(out, x.contiguous(), group=group)
    else:  # gloo has no reduce_scatter: all-reduce then slice
        y = x.contiguous().clone()
        dist.all_reduce(y, group=group)
        r = dist.get_rank(group)
        out.copy_(y[r * out.shape[0]:(r + 1) * out.shape[0]])
    return out


class _ColumnSP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, tp):
        xg = tp.all_gather_rows(x)
        y = _mm(xg, w, bias)
        ctx.save_for_backward(xg, w)
        ctx.tp, ctx.has_bias = tp, bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        xg, w = ctx.saved_tensors
        dy = dy.contiguous()
        dxg = _mm_dgrad(dy, w)
        dx = ctx.tp.reduce_scatter_rows(dxg)
        dw = _mm_wgrad(xg, dy, w)
        return dx, dw, (dy.sum(0) if ctx.has_bias else None), None


class _RowSP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, residual, tp):
        yp = _mm(x, w)
        y = tp.reduce_scatter_rows(yp)
        if bias is not None:
            y = y + bias
        if residual is not None:
            y = y + residual
        ctx.save_for_backward(x, w)
        ctx.tp, ctx.has_bias, ctx.has_res = tp, bias is not None, residual is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dyg = ctx.tp.all_gather_rows(dy.contiguous())
        dx = _mm_dgrad(dyg, w)
        dw = _mm_wgrad(x, dyg, w)
        return dx, dw, (dy.sum(0) if ctx.has_bias else None), (dy if ctx.has_res else None), None


class _CopyToTP(torch.autograd.Function):
    """Megatron `f`: identity forward, all-reduce backward."""

    @staticmethod
    def forward(ctx, x, tp):
        ctx.tp = tp
        return x

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous().clone()
        dist.all_reduce(dy, group=ctx.tp.group)
        return dy, None


class _ReduceFromTP(torch.autograd.Function):
    """Megatron `g`: all-reduce forward, identity backward."""

    @staticmethod
    def forward(ctx, x, tp):
        x = x.contiguous().clone()
        dist.all_reduce(x, group=tp.group)
        return x

    @staticmethod
    def backward(ctx, dy):
        return dy, None


class _ScatterSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, tp):
        ctx.tp = tp
        n = x.shape[0] // tp.world
        return x[tp.rank * n:(tp.rank + 1) * n].contiguous()

    @staticmethod
    def backward(ctx, dy):
        return ctx.tp.all_gather_rows(dy.contiguous()), None


class _VocabParallelCE(torch.autograd.Function):
    """Cross-entropy over vocab-sharded logits: one (max, sum-exp, target-logit) exchange.

    logits: [T, V/tp] local shard (overwritten with its gradient); targets: [T] global ids."""

    @staticmethod
    def forward(ctx, logits, targets, tp, ignore_index):
        T, Vl = logits.shape
        lo = tp.rank * Vl
        if logits.is_cuda and logits.dtype == _lib.HALF and Vl % 8 == 0 and logits.stride(1) == 1 \
                and logits.stride(0) % 8 == 0:
            return _VocabParallelCE._forward_native(ctx, logits, targets.reshape(-1).to(torch.int64).contiguous(),
                                                    tp, ignore_index, lo)
        lf = logits.float()
        local_max = lf.max(-1).values
        gmax = local_max.clone()
        dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=tp.group)
        ex = torch.exp(lf - gmax[:, None])
        stats = torch.zeros(T, 2, dtype=torch.float32, device=logits.device)
        stats[:, 0] = ex.sum(-1)
        mine = (targets >= lo) & (targets < lo + Vl)
        idx = (targets - lo).clamp(0, Vl - 1)
        stats[:, 1] = torch.where(mine, lf.gather(1, idx[:, None]).squeeze(1),
                                  torch.zeros((), device=logits.device))
        dist.all_reduce(stats, group=tp.group)   # sum-exp and target logit in one message
        valid = targets != ignore_index
        nvalid = valid.sum().clamp(min=1).float()
        loss = torch.where(valid, gmax + torch.log(stats[:, 0]) - stats[:, 1],
                           torch.zeros((), device=logits.device))
        # gradient, in place
        p = ex / stats[:, 0:1]
        p[torch.arange(T, device=logits.device)[mine], idx[mine]] -= 1.0
        p = p * valid[:, None]
        logits.copy_(p.to(logits.dtype))
        ctx.save_for_backward(logits, nvalid)
        return loss.sum() / nvalid

    @staticmethod
    def _forward_native(ctx, logits, targets, tp, ignore_index, lo):
        """K9 (csrc/kernels/elementwise.cu): statistics kernel -> one exchange -> loss + in-place gradient."""
        import ctypes

        from ..ops import _lib
        from ..ops._lib import c_float, c_int, c_ll, ptr, stream_ptr

        T, Vl = logits.shape
        dev = logits.device
        if tp.fused is not None:
            stats = tp.fused.vocab_ce_exchange(logits, targets, lo)      # peers stored into our buffer
        else:
            stats = torch.empty(tp.world * T * 3, dtype=torch.float32, device=dev)
            ptrs = (ctypes.c_longlong * tp.world)(*[stats.data_ptr() if r == tp.rank else 0 for r in range(tp.world)])
            _lib.call("b2f_vocab_ce_stats", ptr(logits), ptr(targets), c_ll(T), c_int(Vl), c_ll(logits.stride(0)),
                      c_ll(lo), c_int(tp.rank), c_int(tp.world), ptrs, stream_ptr())
            if tp.world > 1:
                mine = stats[tp.rank * T * 3:(tp.rank + 1) * T * 3].clone()
                dist.all_gather_into_tensor(stats, mine, group=tp.group)
        loss = torch.empty(T, dtype=torch.float32, device=dev)
        _lib.call("b2f_vocab_ce_finish", ptr(logits), ptr(targets), c_ll(T), c_int(Vl), c_ll(logits.stride(0)),
                  c_ll(lo), c_int(tp.world), ptr(stats), c_ll(ignore_index), c_float(1.0), ptr(loss), stream_ptr())
        nvalid = (targets != ignore_index).sum().clamp(min=1).float()
        ctx.save_for_backward(logits, nvalid)
        return loss.sum() / nvalid

    @staticmethod
    def backward(ctx, dloss):
        g, nvalid = ctx.saved_tensors
        return g * (dloss / nvalid).to(g.dtype), None, None, None


class TPContext:
    def __init__(self, group=None, world: int = 1, rank: int = 0, sequence_parallel: bool = True,
                 fused=None):
        self.group, self.world, self.rank = group, world, rank
        self.sequence_parallel = sequence_parallel
        self.fused = fused  # parallel.fused_tp.FusedTP or None

    # -- collectives used by the autograd functions (fused path overrides them)
    def all_gather_rows(self, x):
        return _all_gather_rows(x, self.group, self.world)

    def reduce_scatter_rows(self, x):
        return _reduce_scatter_rows(x, self.group, self.world)

    # -- layer interface ------------------------------------------------------
    def _fused_replicated(self, x, n_allreduced: int) -> bool:
        """Replicated-activation (no sequence parallelism) layers can use K4' when the all-reduced matrix
        [tokens, n] fits its tiling; B2F_FUSED_AR=0 keeps the NCCL all-reduce."""
        return (self.fused is not None and x.is_cuda and x.dim() == 2 and _FUSED_AR
                and self.fused.ar_supported(x.shape[0], n_allreduced))

    def column_linear(self, x, w, bias=None):
        if self.world == 1:
            return G.linear(x, w, bias)
        if self.sequence_parallel:
            if self.fused is not None and x.is_cuda:
                return self.fused.column_linear(x, w, bias)
            return _ColumnSP.apply(x, w, bias, self)
        if self._fused_replicated(x, w.shape[1]):    # dgrad + all-reduce in one kernel (K4')
            return self.fused.column_linear_replicated(x, w, bias)
        return G.linear(_CopyToTP.apply(x, self), w, bias)

    def column_linear_gelu(self, x, w, bias=None):
        if self.world == 1:
            return G.linear_gelu(x, w, bias)
        return E.bias_gelu(self.column_linear(x, w, None), bias)

    def row_linear(self, x, w, bias=None, residual=None):
        if self.world == 1:
            return G.linear(x, w, bias, residual)
        if self.sequence_parallel:
            if self.fused is not None and x.is_cuda:
                return self.fused.row_linear(x, w, bias, residual)
            return _RowSP.apply(x, w, bias, residual, self)
        if self._fused_replicated(x, w.shape[0]):    # GEMM + all-reduce in one kernel (K4')
            return self.fused.row_linear_replicated(x, w, bias, residual)
        y = _ReduceFromTP.apply(G.linear(x, w), self)
        if bias is not None:
            y = y + bias
        return y + residual if residual is not None else y

    def scatter_sequence(self, x):
        if self.world == 1 or not self.sequence_parallel:
            return x
        return _ScatterSeq.apply(x, self)

    def gather_sequence(self, x):
        """Inverse of scatter_sequence (used before non-parallel heads)."""
        if self.world == 1 or not self.sequence_parallel:
            return x
        return _GatherSeq.apply(x, self)

    def lm_head_loss(self, y, weight, targets, ignore_index: int = -100):
        if self.world == 1:
            logits = G.linear(y, weight)
            return E.cross_entropy(logits, targets, ignore_index)
        logits = self.column_linear(y, weight)
        return _VocabParallelCE.apply(logits, targets, self, ignore_index)


class _GatherSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, tp):
        ctx.tp = tp
        return tp.all_gather_rows(x)

    @staticmethod
    def backward(ctx, dy):
        n = dy.shape[0] // ctx.tp.world
        return dy[ctx.tp.rank * n:(ctx.tp.rank + 1) * n].contiguous(), None


SINGLE = TPContext()


def allreduce_replicated_grads(flat_or_params, tp: TPContext) -> None:
    """Sum, across the TP group, the grads of parameters every TP rank holds a copy of and
    feeds with a *sequence shard* (norm weights, row-parallel biases) — Megatron's
    `allreduce_sequence_parallel_gradients`."""
    if tp.world == 1 or not tp.sequence_parallel:
        return
    bufs = [p.main_grad for _, p in flat_or_params.named_parameters()
            if getattr(p, "tp_replicated", False) and getattr(p, "main_grad", None) is not None]
    if not bufs:
        return
    flat = torch.cat([b.reshape(-1) for b in bufs])
    dist.all_reduce(flat, group=tp.group)
    off = 0
    for b in bufs:
        b.copy_(flat[off:off + b.numel()].view_as(b))
        off += b.numel()
