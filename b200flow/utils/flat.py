"""Flat parameter / gradient buffers.

All parameters of a model (or pipeline stage) live in ONE contiguous bf16 buffer
and their gradients in ONE contiguous fp32 ``main_grad`` buffer (each parameter
gets a ``.main_grad`` view that the wgrad GEMM epilogue accumulates into).  This
is what makes the optimizer a single fused launch, the ZeRO-1 partition a pair of
offsets, the DDP bucket a slice, and the Megatron-DeepSpeed checkpoint's flat
optimizer groups trivial to write.

Works on CPU tensors too (used by the gloo tests).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
from torch import nn


@dataclass
class ParamSlot:
    name: str
    offset: int
    numel: int
    shape: tuple


class FlatParams:
    """Re-homes the parameters of `module` into flat storage (in registration order)."""

    def __init__(self, module: nn.Module, *, grad_dtype: torch.dtype = torch.float32,
                 align: int = 128, pad_to_multiple: int = 1, allocator=None, param_filter=None):
        """allocator(numel, dtype) -> (tensor, handle): where the flat buffers live
        (symmetric peer memory for the fused collectives); default = ordinary tensors."""
        params = [(n, p) for n, p in module.named_parameters()
                  if p.requires_grad and (param_filter is None or param_filter(n, p))]
        # de-duplicate tied parameters
        seen, uniq = set(), []
        for n, p in params:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append((n, p))
        assert uniq, "module has no trainable parameters"
        dtype, device = uniq[0][1].dtype, uniq[0][1].device
        self.slots: list[ParamSlot] = []
        off = 0
        for n, p in uniq:
            assert p.dtype == dtype, f"mixed parameter dtypes ({n}: {p.dtype} vs {dtype})"
            self.slots.append(ParamSlot(n, off, p.numel(), tuple(p.shape)))
            off += (p.numel() + align - 1) // align * align
        total = (off + pad_to_multiple - 1) // pad_to_multiple * pad_to_multiple
        self.numel = total
        self.params_handle = self.grads_handle = None
        if allocator is not None:
            self.params, self.params_handle = allocator(total, dtype)
            self.grads, self.grads_handle = allocator(total, grad_dtype)
        else:
            self.params = torch.zeros(total, dtype=dtype, device=device)
            self.grads = torch.zeros(total, dtype=grad_dtype, device=device)
        self.module = module
        self._by_name = dict(uniq)
        for slot, (n, p) in zip(self.slots, uniq):
            view = self.params[slot.offset: slot.offset + slot.numel].view(slot.shape)
            view.copy_(p.data)
            p.data = view
            p.main_grad = self.grads[slot.offset: slot.offset + slot.numel].view(slot.shape)
            p._flat_owner = self   # lets optimizers clear all gradients with one fill

    def named_parameters(self):
        for slot in self.slots:
            yield slot.name, self._by_name[slot.name]

    def zero_grad(self) -> None:
        self.grads.zero_()
        for _, p in self.named_parameters():
            p.grad = None

    def collect_autograd_grads(self) -> None:
        """Fold any gradients autograd left in `.grad` (ops without a main_grad epilogue,
        e.g. embeddings) into the flat fp32 buffer."""
        for _, p in self.named_parameters():
            if p.grad is not None:
                p.main_grad.add_(p.grad.to(p.main_grad.dtype))
                p.grad = None

    def slot(self, name: str) -> ParamSlot:
        for s in self.slots:
            if s.name == name:
                return s
        raise KeyError(name)


def partition_range(numel: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous equal partition [start, end) of a (padded) flat buffer across `world` ranks."""
    assert numel % world == 0, "flat buffer must be padded to a multiple of the partition count"
    per = numel // world
    return rank * per, (rank + 1) * per
