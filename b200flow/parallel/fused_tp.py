"""Fused tensor-parallel linears: K3 all-gather→GEMM, K4 GEMM→reduce-scatter and K4′ GEMM→all-reduce
(csrc/comm/fused_gemm.cu) behind the `TPContext` interface.

Buffer protocol (see the kernel file for the in-kernel part):
* the sequence shard that peers will pull is copied into one of two alternating symmetric
  slots before the K3 launch; a slot is rewritten two K3 calls later, by which time every
  peer has signalled the start of its next K3 and therefore finished reading it;
* partial output tiles land in one of two alternating symmetric receive buffers; a rank can
  only finish call e+1 after every peer produced its call-(e+1) partials, i.e. after the peer
  finished reading call e — so two buffers suffice;
* every flag holds a per-kind call counter (epoch) that only grows.
"""
from __future__ import annotations

import ctypes

import torch

from ..comm.arena import PeerComm
from ..ops import _lib
from ..ops import gemm as G
from ..ops._lib import c_int, c_ll, ptr, stream_ptr


class FusedTP:
    FLAG_WORDS = 1 << 16

    def __init__(self, comm: PeerComm, comm_ctas: int = 16):
        self.comm, self.comm_ctas = comm, comm_ctas
        self.ag_calls = 0
        self.rs_calls = 0
        self._x_slots: dict[int, list] = {}
        self._recv_slots: dict[int, list] = {}
        self.ag_ready = comm.alloc(4 * 64, zero=True)
        self.rs_flags = [comm.alloc(4 * self.FLAG_WORDS, zero=True) for _ in range(2)]
        self.tile_flags = torch.zeros(4096, dtype=torch.int32, device=comm.device)
        self._out_slots: dict[int, list] = {}
        self.ar_done = comm.alloc(4 * 64, zero=True)
        self.ar_counter = torch.zeros(1, dtype=torch.int32, device=comm.device)

    @classmethod
    def create(cls, group, rank, world, device):
        return cls(PeerComm(group, rank, world, device))

    # ------------------------------------------------------------ primitives
    def _slots(self, table, nbytes):
        key = (nbytes + (1 << 20) - 1) >> 20 << 20
        if key not in table:
            table[key] = [self.comm.alloc(key) for _ in range(2)]
        return table[key]

    def ag_gemm(self, x_shard: torch.Tensor, w: torch.Tensor, bias=None, b_mn_major: bool = False):
        """Returns (y[T, N], xg[T, K]) with T = world * rows(x_shard)."""
        comm = self.comm
        ts, k = x_shard.shape
        t = ts * comm.world
        n = w.shape[1] if b_mn_major else w.shape[0]
        self.ag_calls += 1
        slot = self._slots(self._x_slots, ts * k * 2)[self.ag_calls & 1]
        slot.view(_lib.HALF, ts * k).copy_(x_shard.reshape(-1))
        xg = torch.empty((t, k), dtype=_lib.HALF, device=x_shard.device)
        y = torch.empty((t, n), dtype=_lib.HALF, device=x_shard.device)
        assert t // 128 <= self.tile_flags.numel()
        _lib.call("b2f_ag_gemm_bf16", slot.ptr_array, ptr(xg), ptr(w), ptr(y), c_int(t), c_int(n),
                  c_int(k), c_ll(w.stride(0)), c_ll(y.stride(0)), c_int(int(b_mn_major)), ptr(bias),
                  self.ag_ready.ptr_array, ptr(self.tile_flags), c_int(comm.rank), c_int(comm.world),
                  c_int(self.comm_ctas), ctypes.c_uint(self.ag_calls), stream_ptr())
        return y, xg

    def gemm_rs(self, x: torch.Tensor, w: torch.Tensor, bias=None, residual=None,
                b_mn_major: bool = False):
        """Returns y_shard[T / world, N] = reduce_scatter(x @ op(w)) (+bias) (+residual)."""
        comm = self.comm
        t, k = x.shape
        n = w.shape[1] if b_mn_major else w.shape[0]
        ts = t // comm.world
        self.rs_calls += 1
        par = self.rs_calls & 1
        recv = self._slots(self._recv_slots, comm.world * ts * n * 2)[par]
        flags = self.rs_flags[par]
        assert comm.world * (ts // 128) * ((n + 255) // 256) <= self.FLAG_WORDS
        y = torch.empty((ts, n), dtype=_lib.HALF, device=x.device)
        _lib.call("b2f_gemm_rs_bf16", ptr(x), ptr(w), ptr(y), c_int(t), c_int(n), c_int(k),
                  c_ll(x.stride(0)), c_ll(w.stride(0)), c_ll(y.stride(0)), c_int(int(b_mn_major)),
                  ptr(bias), ptr(residual), c_ll(residual.stride(0) if residual is not None else 0),
                  recv.ptr_array, flags.ptr_array, c_int(comm.rank), c_int(comm.world),
                  ctypes.c_uint(self.rs_calls), stream_ptr())
        return y

    def ar_supported(self, t: int, n: int) -> bool:
        return t % (self.comm.world * 128) == 0 and n % 32 == 0

    def gemm_ar(self, x: torch.Tensor, w: torch.Tensor, bias=None, residual=None,
                b_mn_major: bool = False):
        """K4′: returns y[T, N] = all_reduce(x @ op(w)) (+bias) (+residual), complete on every rank.
        Shares K4's receive buffers, flags and call counter; the output lands in one of two alternating
        symmetric slots (peers write their finished rows into it) and is copied out, because callers keep
        it for backward while the slot is reused two calls later."""
        comm = self.comm
        t, k = x.shape
        n = w.shape[1] if b_mn_major else w.shape[0]
        ts = t // comm.world
        assert self.ar_supported(t, n), "gemm_ar: rows per rank must be a multiple of 128, N of 32"
        self.rs_calls += 1
        par = self.rs_calls & 1
        recv = self._slots(self._recv_slots, comm.world * ts * n * 2)[par]
        flags = self.rs_flags[par]
        assert comm.world * (ts // 128) * ((n + 255) // 256) <= self.FLAG_WORDS
        out = self._slots(self._out_slots, t * n * 2)[par]
        res = residual[comm.rank * ts:(comm.rank + 1) * ts] if residual is not None else None
        _lib.call("b2f_gemm_ar_bf16", ptr(x), ptr(w), out.ptr_array, c_int(t), c_int(n), c_int(k),
                  c_ll(x.stride(0)), c_ll(w.stride(0)), c_ll(n), c_int(int(b_mn_major)),
                  ptr(bias), ptr(res), c_ll(res.stride(0) if res is not None else 0),
                  recv.ptr_array, flags.ptr_array, self.ar_done.ptr_array, ptr(self.ar_counter),
                  c_int(comm.rank), c_int(comm.world), ctypes.c_uint(self.rs_calls), stream_ptr())
        return out.view(_lib.HALF, t * n).view(t, n).clone()

    def vocab_ce_exchange(self, logits: torch.Tensor, targets: torch.Tensor, lo: int) -> torch.Tensor:
        """K9 statistics pass with the exchange done by the kernel's stores: every rank writes its
        per-row {max, sum-exp, target logit} into slot `rank` of all peers' [world, T, 3] buffer;
        one flag barrier later each rank holds every shard's statistics.  Returns that buffer."""
        comm = self.comm
        T, vl = logits.shape
        self.ce_calls = getattr(self, "ce_calls", 0) + 1
        buf = self._slots(self.__dict__.setdefault("_ce_slots", {}), comm.world * T * 3 * 4)[self.ce_calls & 1]
        _lib.call("b2f_vocab_ce_stats", ptr(logits), ptr(targets), c_ll(T), c_int(vl), c_ll(logits.stride(0)),
                  c_ll(lo), c_int(comm.rank), c_int(comm.world), buf.ptr_array, stream_ptr())
        if not hasattr(self, "_ce_flags"):
            self._ce_flags = comm.flag_region(3)
        _lib.call("b2f_peer_barrier", self._ce_flags, c_int(comm.rank), c_int(comm.world),
                  ctypes.c_uint(comm.next_epoch()), stream_ptr())
        return buf.view(torch.float32, comm.world * T * 3)

    # ------------------------------------------------------- layer interface
    def column_linear(self, x, w, bias=None):
        return _FusedColumn.apply(x, w, bias, self)

    def row_linear(self, x, w, bias=None, residual=None):
        return _FusedRow.apply(x, w, bias, residual, self)

    # the same two layers with replicated activations (no sequence parallelism)
    def column_linear_replicated(self, x, w, bias=None):
        return _FusedColumnAR.apply(x, w, bias, self)

    def row_linear_replicated(self, x, w, bias=None, residual=None):
        return _FusedRowAR.apply(x, w, bias, residual, self)


class _FusedColumn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, ftp):
        y, xg = ftp.ag_gemm(x.contiguous(), w, bias)
        ctx.save_for_backward(xg, w)
        ctx.ftp, ctx.has_bias = ftp, bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        xg, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = ctx.ftp.gemm_rs(dy, w, b_mn_major=True)          # dgrad fused with its reduce-scatter
        dw = G._wgrad(xg, dy, w)
        return dx, dw, (dy.sum(0) if ctx.has_bias else None), None


class _FusedRow(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, residual, ftp):
        x = x.contiguous()
        y = ftp.gemm_rs(x, w, bias, residual.contiguous() if residual is not None else None)
        ctx.save_for_backward(x, w)
        ctx.ftp, ctx.has_bias, ctx.has_res = ftp, bias is not None, residual is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx, dyg = ctx.ftp.ag_gemm(dy, w, b_mn_major=True)      # all-gather fused with dgrad
        dw = G._wgrad(x, dyg, w)
        return dx, dw, (dy.sum(0) if ctx.has_bias else None), (dy if ctx.has_res else None), None


class _FusedColumnAR(torch.autograd.Function):
    """Megatron `f` + column-parallel linear: plain GEMM forward, dgrad fused with its all-reduce (K4′)."""

    @staticmethod
    def forward(ctx, x, w, bias, ftp):
        x = x.contiguous()
        ctx.save_for_backward(x, w)
        ctx.ftp, ctx.has_bias = ftp, bias is not None
        return G.gemm(x, w, bias=bias)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = ctx.ftp.gemm_ar(dy, w, b_mn_major=True)
        dw = G._wgrad(x, dy, w)
        return dx, dw, (dy.sum(0) if ctx.has_bias else None), None


class _FusedRowAR(torch.autograd.Function):
    """Row-parallel linear + Megatron `g`: GEMM fused with the all-reduce (K4′); backward needs no collective."""

    @staticmethod
    def forward(ctx, x, w, bias, residual, ftp):
        x = x.contiguous()
        y = ftp.gemm_ar(x, w, bias, residual.contiguous() if residual is not None else None)
        ctx.save_for_backward(x, w)
        ctx.has_bias, ctx.has_res = bias is not None, residual is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = G.gemm(dy, w, b_mn_major=True)
        dw = G._wgrad(x, dy, w)
        return dx, dw, (dy.sum(0) if ctx.has_bias else None), (dy if ctx.has_res else None), None
