"""Fused data-parallel paths over peer memory: K2 (ZeRO-1) and K1 (DDP buckets).

`FusedZero1.step` replaces the three NCCL phases of a ZeRO-1 step with two kernels
(csrc/comm/collectives.cu):  K2a pulls every peer's fp32 gradient for the owned shard
over NVLink, reduces it and produces the shard's sum of squares; after the (scalar) norm
all-reduce, K2b clips, runs AdamW on the fp32 master shard and stores the bf16 result
straight into every rank's parameter buffer.  Gradients and parameters therefore live in
symmetric memory (`FlatParams(allocator=comm.allocator())`).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..comm.arena import PeerComm
from ..ops import _lib
from ..ops import optim as O
from ..ops._lib import c_float, c_int, c_ll, ptr, stream_ptr
import ctypes


class FusedZero1:
    def __init__(self, comm: PeerComm, grads_buf, params_buf):
        self.comm, self.grads_buf, self.params_buf = comm, grads_buf, params_buf
        self.blocks = min(128, comm.max_blocks())      # nothing to overlap with at the end of a step: every SM helps
        self.flags = comm.flag_region(1)
        dev = comm.device
        self.partial = torch.zeros(self.blocks, dtype=torch.float32, device=dev)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.reduced = None
        import os

        # in-switch reduction / replication pays once more than one peer feeds a shard: at two GPUs plain peer loads and
        # stores were faster (13.6 vs 15.7 ms at 0.88 G elements, profiles/r2_call8_*), so it defaults on from three up
        self.use_multicast = os.environ.get("B2F_ZERO1_MULTIMEM", "1" if comm.world > 2 else "0") == "1"

    @classmethod
    def create(cls, group, rank, world, device):
        """Returns (comm, allocator); the engine builds FlatParams with the allocator and then
        calls `bind(flat)`."""
        return PeerComm(group, rank, world, device)

    def step(self, opt, div: float):
        comm, flat, cfg = self.comm, opt.flat, opt.cfg
        n = opt.hi - opt.lo
        if self.reduced is None:
            self.reduced = torch.empty(n, dtype=torch.float32, device=comm.device)
        gmc = int(getattr(self.grads_buf, "multicast_ptr", 0) or 0) if self.use_multicast else 0
        if gmc:     # the switch sums the shard (multimem.ld_reduce): (world - 1) x fewer bytes into this GPU
            _lib.call("b2f_zero1_reduce_mc", self.grads_buf.ptr_array, c_ll(gmc), self.flags, c_int(comm.rank),
                      c_int(comm.world), c_ll(opt.lo), c_ll(n), ptr(self.reduced), ptr(self.partial),
                      ptr(self.sumsq), c_int(self.blocks), ctypes.c_uint(comm.next_epoch()), stream_ptr())
        else:
            _lib.call("b2f_zero1_reduce", self.grads_buf.ptr_array, self.flags, c_int(comm.rank),
                      c_int(comm.world), c_ll(opt.lo), c_ll(n), ptr(self.reduced), ptr(self.partial),
                      ptr(self.sumsq), c_int(self.blocks), ctypes.c_uint(comm.next_epoch()), stream_ptr())
        # global norm: shard sums over DP, then over the model-parallel groups
        dist.all_reduce(self.sumsq, group=opt.dp_group)
        for grp in opt.norm_groups:
            if grp is not None:
                dist.all_reduce(self.sumsq, group=grp)
        inv = 1.0 / div
        self.sumsq.mul_(inv * inv)
        coef, nrm = O.clip_coef(self.sumsq, cfg.clip_grad, extra_scale=inv)
        pmc = int(getattr(self.params_buf, "multicast_ptr", 0) or 0) if self.use_multicast else 0
        _lib.call("b2f_zero1_adam_gather_mc", ptr(self.reduced), ptr(opt.master), ptr(opt.exp_avg),
                  ptr(opt.exp_avg_sq), self.params_buf.ptr_array, c_ll(pmc), self.flags, c_int(comm.rank),
                  c_int(comm.world), c_ll(opt.lo), c_ll(n), c_float(opt.lr), c_float(cfg.beta1),
                  c_float(cfg.beta2), c_float(cfg.eps), c_float(cfg.weight_decay),
                  c_int(opt.step_count), ptr(coef), ptr(None), c_int(self.blocks),
                  ctypes.c_uint(comm.next_epoch()), stream_ptr())
        return nrm


class FusedAllReduce:
    """K1: in-place averaged all-reduce of slices of a symmetric fp32 gradient buffer."""

    def __init__(self, comm: PeerComm, grads_buf, blocks: int = 32, use_multicast: bool = True,
                 wire_bf16: bool | None = None):
        import os

        self.comm, self.grads_buf = comm, grads_buf
        self.blocks = min(blocks, comm.max_blocks())
        self.flags = comm.flag_region(2)
        self.use_multicast = use_multicast
        self.mc = grads_buf.multicast_ptr if use_multicast else 0
        # bf16 on the wire (half the NVLink bytes; the gradients stay fp32 in the flat buffer): B2F_DDP_WIRE=fp32 keeps
        # the all-fp32 kernel
        self.wire_bf16 = (os.environ.get("B2F_DDP_WIRE", "bf16") == "bf16") if wire_bf16 is None else wire_bf16
        self._stage, self._stage_n = None, 0

    def _staging(self, n: int):
        """Symmetric bf16 staging of at least n elements (every rank asks for the same sizes in the same order)."""
        if n > self._stage_n:
            self._stage = self.comm.alloc(max(n, 1 << 20) * 2)
            self._stage_n = max(n, 1 << 20)
        return self._stage

    def __call__(self, off: int, n: int, scale: float, bf16_out: torch.Tensor | None = None):
        comm = self.comm
        if self.wire_bf16 and bf16_out is None and n % 8 == 0 and off % 4 == 0:
            st = self._staging(n)
            smc = int(st.multicast_ptr or 0) if self.use_multicast else 0
            mine = self.grads_buf.ptrs[comm.rank]
            _lib.call("b2f_allreduce_bf16wire", ctypes.c_void_p(mine), st.ptr_array, c_ll(smc), self.flags,
                      c_int(comm.rank), c_int(comm.world), c_ll(off), c_ll(n), c_float(scale), c_int(self.blocks),
                      ctypes.c_uint(comm.next_epoch(2)), stream_ptr())
            return
        _lib.call("b2f_allreduce_f32", self.grads_buf.ptr_array, c_ll(self.mc), self.flags,
                  c_int(comm.rank), c_int(comm.world), c_ll(off), c_ll(n), c_float(scale),
                  ptr(bf16_out), c_int(self.blocks), ctypes.c_uint(comm.next_epoch(2)), stream_ptr())
