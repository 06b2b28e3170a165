"""K5 behind the MoE layer: the expert FFN of an expert-parallel layer as two fused kernels per direction
(csrc/comm/fused_moe.cu) — dispatch + block-scaled fp8 GEMM ("up"), block-scaled fp8 GEMM + combine ("down").

Forward of one layer on every rank (`T` tokens, `E` experts over `world` ranks, capacity `C` per (rank, expert)):

    token_of            slot -> token map of the routing decision            (tiny kernel)
    UP(x, W1)           tokens leave as e4m3 + UE8M0 scales into the owners' buffers, expert GEMM 1 starts tile by tile
                        as they land; epilogue: + b1, keep pre-activation, GELU, quantise for GEMM 2
    DOWN(h, W2)         expert GEMM 2; epilogue: + b2, rows go home to the rank the token came from
    wait, combine       the source applies the gate and un-permutes                   (local kernels)

Backward is the same pair with the transposed weights: UP(dy * gate, W2^T) with the epilogue "x gelu'(pre)", then
DOWN(dpre, W1^T) bringing dx rows home; the weight gradients are bf16 tcgen05 GEMMs with fp32 accumulation over the
dequantised operands (the scale blocks run along the hidden dimension, the weight gradient contracts over tokens).
The gate gradient needs no second trip: y_t = gate_t * row_t, so d gate_t = <dy_t, y_t> / gate_t.

Buffer reuse across calls.  The symmetric token buffers, the rows-home buffers and the completion counters exist twice
and alternate with a call counter that all ranks advance in lock step; flags and token counts carry that counter.
A rank can be at most one call ahead of a peer: it cannot finish call e (its rows only come home from a peer's DOWN
kernel of call e, which runs after that peer's UP kernel of call e has consumed the tokens of call e) before every peer
has at least launched call e, so when it writes parity (e+1) & 1 everyone is done reading the data of call e-1.  The
counters of the other parity are cleared on the stream right before a call's kernels, i.e. before this rank's own
tokens of the call leave, without which no peer can complete the call.

Reference capability: DeepSpeed-MoE's `MOELayer` (einsum dispatch -> `all_to_all_single` -> experts -> `all_to_all_single`
-> einsum combine) as pinned by containers/megatron-deepspeed/Dockerfile:12-13 and switched on by the `--num-experts /
--moe-expert-parallel-size` flags of Megatron-DeepSpeed's `pretrain_gpt.py` (the entry point of
examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:74-85); SURVEY.md §2F "EP", §2H C13-C15, §8 K5.
"""
from __future__ import annotations

import ctypes

import torch

from ..ops import _lib
from ..ops import elementwise as E_
from ..ops import gemm as G
from ..ops import gemm_mx as MX
from ..ops._lib import c_int, ptr, stream_ptr


def supported(hidden: int, ffn: int, capacity: int, tokens: int) -> bool:
    return capacity % 128 == 0 and hidden % 512 == 0 and ffn % 512 == 0 and tokens > 0


class _Local:
    """Stand-in for a symmetric buffer when the expert-parallel group is one GPU."""

    def __init__(self, nbytes, device):
        self.tensor = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        self.ptrs = [self.tensor.data_ptr()]
        self.ptr_array = (ctypes.c_longlong * 1)(*self.ptrs)

    def view(self, dtype, numel=None, offset_bytes=0):
        t = self.tensor[offset_bytes:].view(dtype)
        return t if numel is None else t[:numel]


class FusedMoE:
    """Buffers and launches of K5 for one expert-parallel group (shared by all MoE layers of the model)."""

    def __init__(self, ep):
        self.ep = ep
        self.world, self.rank = max(ep.world, 1), ep.rank
        self.calls = 0
        self._bufs: dict = {}
        self._wq: dict = {}
        # the optimizer updates parameters through raw pointers, so tensor versions do not tell when a weight changed:
        # an engine that calls `invalidate()` after every optimizer step may switch the cache on
        self.cache_weights = False

    def invalidate(self):
        self._wq.clear()

    def _alloc(self, key, nbytes, device):
        k = (key, (nbytes + (1 << 20) - 1) >> 20)
        if k not in self._bufs:
            make = (lambda: self.ep.comm.alloc(k[1] << 20, zero=True)) if self.ep.comm is not None else \
                (lambda: _Local(k[1] << 20, device))
            self._bufs[k] = [make() for _ in range(2)]
        return self._bufs[k]

    def quantized_weight(self, w: torch.Tensor, transpose: bool):
        """MX copy of an expert weight [El, N, K] (or of its transpose), refreshed when the weight changes."""
        key = (id(w), transpose)
        hit = self._wq.get(key) if self.cache_weights else None
        if hit is None:
            hit = MX.quantize_mx_t(w.detach()) if transpose else MX.quantize_mx(w.detach())
            if self.cache_weights:
                self._wq[key] = hit
        return hit

    # ------------------------------------------------------------------ one direction
    def run(self, x, token_of, counts, E, C, w_up, bias_up, w_down, bias_down, pre=None):
        """x [T, H] bf16 tokens of this rank -> (rows_home [E, C, Hout] bf16 view, keep [El, R, F] bf16, xhat [El, R, H] bf16).
        `pre` given: backward epilogue (x gelu'(pre)); w_up: MxTensor [El, F, H]; w_down: MxTensor [El, Hout, F]."""
        dev = x.device
        T, H = x.shape
        W, El = self.world, E // self.world
        R = W * C
        F = w_up.rows
        Hout = w_down.rows
        self.calls += 1
        par, epoch = self.calls & 1, self.calls
        xq = self._alloc("xq", El * R * H, dev)[par]
        xsf = self._alloc("xsf", El * MX.sf_bytes(R, H), dev)[par]
        flags = self._alloc("flags", El * (R // 128) * 4, dev)[0]
        cnt = self._alloc("cnt", W * El * 4, dev)[par]     # a peer one call ahead already writes the other parity
        rows = self._alloc("rows", E * C * Hout * 2, dev)[par]
        done = self._alloc("done", 256, dev)
        done[1 - par].view(torch.int32, 1).zero_()        # next call's counter; see the module docstring
        keep = torch.empty((El, R, F), dtype=_lib.HALF, device=dev)
        hq = torch.empty((El, R, F), dtype=torch.uint8, device=dev)
        hsf = torch.empty(El * MX.sf_bytes(R, F), dtype=torch.uint8, device=dev)
        _lib.call("b2f_moe_up", ptr(x), ptr(token_of), ptr(counts), c_int(T), c_int(H), c_int(E), c_int(C),
                  xq.ptr_array, xsf.ptr_array, flags.ptr_array, cnt.ptr_array, ptr(w_up.q), ptr(w_up.sf), c_int(F),
                  ptr(bias_up), ptr(keep), ptr(pre), ptr(hq), ptr(hsf), c_int(int(pre is not None)),
                  c_int(self.rank), c_int(W), ctypes.c_uint(epoch), stream_ptr())
        my_cnt = cnt.view(torch.int32, W * El)
        # the operands the bf16 weight-gradient GEMMs need, as whole-buffer tensors with zeros where no token sits
        xhat = MX.dequantize_mx(MX.MxTensor(xq.view(torch.uint8, El * R * H).view(El, R, H),
                                            xsf.view(torch.uint8, El * MX.sf_bytes(R, H))))
        _lib.call("b2f_moe_down", ptr(hq), ptr(hsf), ptr(w_down.q), ptr(w_down.sf), c_int(F), c_int(Hout),
                  ptr(bias_down), rows.ptr_array, done[par].ptr_array, ptr(my_cnt), c_int(E), c_int(C),
                  c_int(self.rank), c_int(W), ctypes.c_uint(epoch), stream_ptr())
        for buf, n in ((xhat, H), (keep, F)):
            _lib.call("b2f_moe_zero_inactive", ptr(buf), ptr(my_cnt), c_int(El), c_int(W), c_int(C), c_int(n), stream_ptr())
        _lib.call("b2f_moe_wait_rows", ptr(done[par].view(torch.int32, 1)), ptr(counts), c_int(E), c_int(Hout), stream_ptr())
        return rows.view(_lib.HALF, E * C * Hout).view(E * C, Hout), keep, xhat


def token_map(expert, slot, E, C):
    T = expert.shape[0]
    token_of = torch.full((E * C,), -1, dtype=torch.int32, device=expert.device)
    _lib.call("b2f_moe_token_of", ptr(expert), ptr(slot), c_int(T), c_int(C), ptr(token_of), stream_ptr())
    return token_of


class _FusedExperts(torch.autograd.Function):
    """y[t] = gate[t] * FFN_{expert[t]}(x[t]) over the expert-parallel group; dropped tokens give zero rows."""

    @staticmethod
    def forward(ctx, x, gate, w1, b1, w2, b2, expert, slot, counts, E, C, fm: FusedMoE):
        from .moe import _combine_local

        x = x.contiguous()
        token_of = token_map(expert, slot, E, C)
        counts = counts.to(torch.int32).contiguous()
        rows, pre, xhat = fm.run(x, token_of, counts, E, C, fm.quantized_weight(w1, False), b1,
                                 fm.quantized_weight(w2, False), b2)
        y = _combine_local(rows, expert, slot, gate, C)
        ctx.save_for_backward(gate, expert, slot, counts, token_of, pre, xhat, y, w1, w2)
        ctx.cfg = (E, C, fm, b1 is not None, b2 is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        from .moe import _combine_local

        gate, expert, slot, counts, token_of, pre, xhat, y, w1, w2 = ctx.saved_tensors
        E, C, fm, has_b1, has_b2 = ctx.cfg
        dy = dy.contiguous()
        kept = expert >= 0
        # y_t = gate_t * row_t  =>  d gate_t = <dy_t, row_t> = <dy_t, y_t> / gate_t
        dgate = torch.where(kept, (dy.float() * y.float()).sum(-1) / gate.clamp(min=1e-20), torch.zeros_like(gate))
        dyg = (dy * gate[:, None].to(dy.dtype)).contiguous()
        rows, dpre, dyhat = fm.run(dyg, token_of, counts, E, C, fm.quantized_weight(w2, True), None,
                                   fm.quantized_weight(w1, True), None, pre=pre)
        ones = torch.ones_like(gate)
        dx = _combine_local(rows, expert, slot, ones, C)
        # weight gradients: bf16 GEMMs, fp32 accumulation, over all rows of each local expert (inactive rows are zero)
        El, R, F = pre.shape
        h = E_.gelu_fwd_raw(pre.view(El * R, F), None).view(El, R, F)
        dw1 = dw2 = db1 = db2 = None
        for e in range(El):
            r2 = G._wgrad(h[e], dyhat[e], _ExpertSlice(w2, e))        # dW2[e] [H, F] += dy_rows^T h
            r1 = G._wgrad(xhat[e], dpre[e], _ExpertSlice(w1, e))      # dW1[e] [F, H] += dpre^T xhat
            if r2 is not None:
                dw2 = torch.zeros_like(w2) if dw2 is None else dw2
                dw2[e] = r2
            if r1 is not None:
                dw1 = torch.zeros_like(w1) if dw1 is None else dw1
                dw1[e] = r1
        for w in (w1, w2):
            cb = getattr(w, "grad_ready", None)      # DDP bucket bookkeeping, once per stacked parameter
            if cb is not None and getattr(w, "main_grad", None) is not None:
                cb()
        if has_b1:
            db1 = torch.sum(dpre, dim=1, dtype=torch.float32).to(dpre.dtype)
        if has_b2:
            db2 = torch.sum(dyhat, dim=1, dtype=torch.float32).to(dyhat.dtype)
        return dx, dgate, dw1, db1, dw2, db2, None, None, None, None, None, None


class _ExpertSlice:
    """The view the wgrad helper needs of one expert of a stacked [El, ...] parameter: its slice of `main_grad`."""

    def __init__(self, w, e):
        mg = getattr(w, "main_grad", None)
        self.main_grad = mg[e] if mg is not None else None
        self.grad_ready = None


def fused_experts(x, gate, w1, b1, w2, b2, expert, slot, counts, E, C, fm):
    return _FusedExperts.apply(x, gate, w1, b1, w2, b2, expert, slot, counts, E, C, fm)
