"""Symmetric peer-memory arena over NVLink 5 / NVSwitch.

A `PeerComm` belongs to one process group.  It owns
  * symmetric buffers: the same-sized allocation on every rank, mapped into every
    peer's address space, optionally with an NVLS multicast mapping;
  * a flag region (epoch counters) for the in-kernel cross-GPU barriers;
  * the epoch counter itself (monotonic, identical on all ranks because every rank
    issues the same sequence of collective kernels).

Allocation and the handle exchange go through `torch.distributed._symmetric_memory`
(cuMem + fabric/fd handle exchange over the c10d store, NVLS multicast object when the
switch supports it) — this is plumbing; every byte that crosses NVLink on a hot path is
moved by b200flow's own kernels (csrc/comm/*.cu) through these pointers.

Stands in for NCCL's communicator bootstrap used by the reference
(`--distributed-backend nccl`, examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:82;
SURVEY.md §2G, §5.8).
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist


class SymmetricBuffer:
    def __init__(self, tensor: torch.Tensor, handle):
        self.tensor = tensor
        self.handle = handle
        self.ptrs = [int(p) for p in handle.buffer_ptrs]
        try:
            self.multicast_ptr = int(handle.multicast_ptr) if handle.has_multicast_support(
                tensor.device.type, tensor.device.index) else 0
        except Exception:  # pragma: no cover - API drift between torch versions
            self.multicast_ptr = int(getattr(handle, "multicast_ptr", 0) or 0)
        self._ptr_array = (ctypes.c_longlong * len(self.ptrs))(*self.ptrs)

    @property
    def ptr_array(self):
        return self._ptr_array

    def view(self, dtype: torch.dtype, numel: int | None = None, offset_bytes: int = 0) -> torch.Tensor:
        raw = self.tensor.view(torch.uint8)[offset_bytes:]
        t = raw.view(dtype)
        return t if numel is None else t[:numel]


class PeerComm:
    MAX_SLOTS = 256

    def __init__(self, group, rank: int, world: int, device: torch.device):
        self.group, self.rank, self.world, self.device = group, rank, world, device
        self.epoch = 1
        self._flags = self.alloc(self.MAX_SLOTS * max(world, 1) * 4 * 8, zero=True)
        self.blocks = None

    # ------------------------------------------------------------------ alloc
    def alloc(self, nbytes: int, zero: bool = False) -> SymmetricBuffer:
        import torch.distributed._symmetric_memory as symm

        nbytes = (nbytes + 255) // 256 * 256
        t = symm.empty(nbytes, dtype=torch.uint8, device=self.device)
        if zero:
            t.zero_()
        hdl = symm.rendezvous(t, self.group)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        return SymmetricBuffer(t, hdl)

    def allocator(self):
        """Returns f(numel, dtype) -> (tensor, SymmetricBuffer) for FlatParams."""
        def f(numel: int, dtype: torch.dtype):
            buf = self.alloc(numel * torch.empty((), dtype=dtype).element_size(), zero=True)
            return buf.view(dtype, numel), buf
        return f

    # ------------------------------------------------------------------ flags
    @property
    def flags(self) -> SymmetricBuffer:
        return self._flags

    def flag_region(self, index: int) -> ctypes.Array:
        """Flag pointers offset to a private region (each kernel family keeps its own slots)."""
        off = index * self.MAX_SLOTS * self.world * 4
        return (ctypes.c_longlong * self.world)(*[p + off for p in self._flags.ptrs])

    def next_epoch(self, count: int = 1) -> int:
        e = self.epoch
        self.epoch += count
        return e

    def max_blocks(self) -> int:
        if self.blocks is None:
            from ..ops import _lib

            self.blocks = min(_lib.call_int("b2f_comm_max_blocks"), self.MAX_SLOTS)
        return self.blocks
