"""K6 / K11 — fused AdamW, l2-norm, clip and scale wrappers (csrc/kernels/optim.cu)."""
from __future__ import annotations

import torch

from . import _lib
from ._lib import c_float, c_int, c_ll, ptr, stream_ptr


def adamw_flat_(master: torch.Tensor, m: torch.Tensor, v: torch.Tensor, grad: torch.Tensor,
                param_out: torch.Tensor | None, *, lr: float, beta1: float, beta2: float,
                eps: float, weight_decay: float, step: int,
                grad_scale: torch.Tensor | None = None, skip: torch.Tensor | None = None) -> None:
    """One AdamW step over a flat fp32 master buffer; writes the bf16 working copy too."""
    _lib.require_cuda(master, grad)
    assert master.dtype == torch.float32 and m.dtype == torch.float32 and v.dtype == torch.float32
    assert grad.dtype in (_lib.HALF, torch.float32) and grad.numel() == master.numel()
    assert master.is_contiguous() and grad.is_contiguous()
    if param_out is not None:
        assert param_out.dtype == _lib.HALF and param_out.numel() == master.numel()
    _lib.call("b2f_adamw_flat", ptr(master), ptr(m), ptr(v), ptr(grad),
              c_int(int(grad.dtype == _lib.HALF)), ptr(param_out), c_ll(master.numel()),
              c_float(lr), c_float(beta1), c_float(beta2), c_float(eps), c_float(weight_decay),
              c_int(step), ptr(grad_scale), ptr(skip), stream_ptr())


class MultiTensorTable:
    """Device-side segment table for lists of separately allocated tensors (amp_C-style)."""

    def __init__(self, masters, ms, vs, grads, params_out):
        chunk = _lib.call_int("b2f_multi_tensor_chunk")
        rows, first = [], 0
        for i, mst in enumerate(masters):
            n = mst.numel()
            # the kernel reads raw pointers as fp32 master / moments of n elements each
            for t in (mst, ms[i], vs[i]):
                assert t.dtype == torch.float32 and t.numel() == n and t.is_contiguous(), \
                    f"AdamW state must be contiguous float32 (got {t.dtype}, {t.numel()} vs {n})"
            assert grads[i].dtype == grads[0].dtype and grads[i].numel() == n
            po = params_out[i].data_ptr() if params_out is not None and params_out[i] is not None else 0
            rows.append([mst.data_ptr(), ms[i].data_ptr(), vs[i].data_ptr(), grads[i].data_ptr(),
                         po, n, first])
            first += (n + chunk - 1) // chunk
        self.total_chunks = first
        self.nsegs = len(rows)
        self.grad_bf16 = grads[0].dtype == _lib.HALF
        self.table = torch.tensor(rows, dtype=torch.int64, device=masters[0].device)
        self._keep = (masters, ms, vs, grads, params_out)


def adamw_multi_(table: MultiTensorTable, *, lr, beta1, beta2, eps, weight_decay, step,
                 grad_scale=None, skip=None) -> None:
    _lib.call("b2f_adamw_multi", ptr(table.table), c_int(table.nsegs), c_ll(table.total_chunks),
              c_int(int(table.grad_bf16)), c_float(lr), c_float(beta1), c_float(beta2), c_float(eps),
              c_float(weight_decay), c_int(step), ptr(grad_scale), ptr(skip), stream_ptr())


def sumsq(x: torch.Tensor, out: torch.Tensor | None = None, overflow: torch.Tensor | None = None,
          accumulate: bool = False, scratch: torch.Tensor | None = None) -> torch.Tensor:
    """out[0] (+)= sum(x^2) in fp32 (deterministic two-stage reduction)."""
    _lib.require_cuda(x)
    assert x.is_contiguous() and x.dtype in (_lib.HALF, torch.float32)
    n = x.numel()
    nparts = _lib.call_int("b2f_sumsq_num_partials", c_ll(n))
    if scratch is None:
        scratch = torch.empty(nparts, dtype=torch.float32, device=x.device)
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=x.device)
        accumulate = False
    _lib.call("b2f_sumsq", ptr(x), c_ll(n), c_int(int(x.dtype == torch.float32)), ptr(scratch),
              ptr(out), ptr(overflow), c_int(int(accumulate)), stream_ptr())
    return out


def clip_coef(sumsq_t: torch.Tensor, max_norm: float, extra_scale: float = 1.0):
    """Returns (coef, norm) device scalars: coef = min(1, max_norm/(norm+1e-6)) * extra_scale."""
    coef = torch.empty(1, dtype=torch.float32, device=sumsq_t.device)
    nrm = torch.empty(1, dtype=torch.float32, device=sumsq_t.device)
    _lib.call("b2f_clip_coef", ptr(sumsq_t), c_float(max_norm), c_float(extra_scale), ptr(coef),
              ptr(nrm), stream_ptr())
    return coef, nrm


def scale_(x: torch.Tensor, scale: float = 1.0, scale_ptr: torch.Tensor | None = None) -> None:
    assert x.is_contiguous() and x.dtype in (_lib.HALF, torch.float32)
    _lib.call("b2f_scale", ptr(x), c_ll(x.numel()), c_int(int(x.dtype == torch.float32)),
              ptr(scale_ptr), c_float(scale), stream_ptr())
