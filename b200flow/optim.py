"""torch.optim-shaped front-ends over the fused optimizer kernels (K11) and LR schedules.

`AdamW` keeps fp32 master weights + moments for (bf16 or fp32) parameters and updates every
tensor of every group in ONE multi-tensor launch (`b2f_adamw_multi`); gradients are taken
from `p.main_grad` (flat fp32 buffer, when the model was flattened) or `p.grad`.  On CPU it
runs the fp32 oracle.  Stands in for apex `FusedAdam` / `torch.optim.AdamW(fused=True)`
(SURVEY.md §2H C23).
"""
from __future__ import annotations

import math

import torch

from .ops import _lib
from .ops import optim as O
from .ops import ref


class AdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._tables = {}
        self._step = 0
        self.grad_scale = None      # optional device scalar multiplied into every gradient
        self.skip = None            # optional device int flag: skip the step (fp16 overflow)

    def _grad(self, p):
        g = getattr(p, "main_grad", None)
        return g if g is not None else p.grad

    def _init_state(self, p):
        st = self.state[p]
        if "master" not in st:
            st["master"] = p.detach().float().clone()
            st["exp_avg"] = torch.zeros_like(st["master"])
            st["exp_avg_sq"] = torch.zeros_like(st["master"])
        return st

    @torch.no_grad()
    def step(self, closure=None):
        self._step += 1
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if self._grad(p) is not None]
            if not ps:
                continue
            b1, b2 = group["betas"]
            if ps[0].is_cuda:
                key = (gi, tuple(id(p) for p in ps), tuple(self._grad(p).data_ptr() for p in ps))
                tab = self._tables.get(key)
                if tab is None:
                    sts = [self._init_state(p) for p in ps]
                    outs = [p.data if p.dtype == _lib.HALF else None for p in ps]
                    tab = O.MultiTensorTable([s["master"] for s in sts], [s["exp_avg"] for s in sts],
                                             [s["exp_avg_sq"] for s in sts],
                                             [self._grad(p).contiguous() for p in ps], outs)
                    self._tables = {key: tab} if len(self._tables) > 8 else {**self._tables, key: tab}
                O.adamw_multi_(tab, lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"],
                               weight_decay=group["weight_decay"], step=self._step,
                               grad_scale=self.grad_scale, skip=self.skip)
                for p in ps:
                    if p.dtype != _lib.HALF:   # fp32 params: master is the truth
                        p.data.copy_(self.state[p]["master"])
            else:
                for p in ps:
                    st = self._init_state(p)
                    g = self._grad(p).float()
                    if self.grad_scale is not None:
                        g = g * float(self.grad_scale)
                    ref.adamw_step(st["master"], st["exp_avg"], st["exp_avg_sq"], g, lr=group["lr"],
                                   beta1=b1, beta2=b2, eps=group["eps"],
                                   weight_decay=group["weight_decay"], step=self._step)
                    p.data.copy_(st["master"].to(p.dtype))

    def state_dict(self):
        sd = super().state_dict()
        sd["b2f_step"] = self._step            # bias correction continues where it stopped
        return sd

    def load_state_dict(self, sd):
        sd = dict(sd)
        self._step = int(sd.pop("b2f_step", self._step))
        super().load_state_dict(sd)
        # torch's loader casts floating-point state to the parameter's dtype; with bf16 parameters that
        # would turn the fp32 master / moments into bf16 while the kernel reads them as fp32.  Restore
        # them from the checkpoint's own tensors, as float32 on the parameter's device.
        saved = sd.get("state", {})
        ids = [i for g in sd.get("param_groups", []) for i in g["params"]]
        params = [p for g in self.param_groups for p in g["params"]]
        for i, p in zip(ids, params):
            src = saved.get(i)
            if src is None:
                continue
            st = self.state[p]
            for k in ("master", "exp_avg", "exp_avg_sq"):
                if k in src:
                    st[k] = src[k].detach().to(device=p.device, dtype=torch.float32).clone()
        self._tables = {}                      # the device tables point at the tensors just replaced

    def zero_grad(self, set_to_none: bool = True):
        owners = {}
        for group in self.param_groups:
            for p in group["params"]:
                p.grad = None
                owner = getattr(p, "_flat_owner", None)
                if owner is not None:
                    owners[id(owner)] = owner      # flat storage: one fill for every gradient
                    continue
                mg = getattr(p, "main_grad", None)
                if mg is not None:
                    mg.zero_()
        for owner in owners.values():
            owner.grads.zero_()


def get_linear_schedule_with_warmup(optimizer, num_warmup_steps: int, num_training_steps: int):
    """HF `get_linear_schedule_with_warmup` (used by the Accelerate example)."""
    def f(step):
        if step < num_warmup_steps:
            return step / max(1, num_warmup_steps)
        return max(0.0, (num_training_steps - step) / max(1, num_training_steps - num_warmup_steps))
    return torch.optim.lr_scheduler.LambdaLR(optimizer, f)


def get_cosine_schedule(optimizer, num_warmup_steps: int, num_training_steps: int, min_ratio: float = 0.0):
    def f(step):
        if step < num_warmup_steps:
            return step / max(1, num_warmup_steps)
        t = (step - num_warmup_steps) / max(1, num_training_steps - num_warmup_steps)
        return min_ratio + (1 - min_ratio) * 0.5 * (1 + math.cos(math.pi * min(1.0, t)))
    return torch.optim.lr_scheduler.LambdaLR(optimizer, f)


class FlatMomentumSGD:
    """Momentum SGD over a `FlatParams` buffer (one fused launch): the optimizer of the
    reference's Mask R-CNN workloads (tensorpack `MomentumOptimizer(0.9)`)."""

    def __init__(self, flat, lr: float = 0.01, momentum: float = 0.9, weight_decay: float = 1e-4):
        self.flat, self.lr, self.momentum, self.weight_decay = flat, lr, momentum, weight_decay
        self.master = flat.params.float().clone()
        self.buf = torch.zeros_like(self.master)
        self.grad_scale = None
        self.param_groups = [{"lr": lr}]

    @torch.no_grad()
    def step(self):
        from .ops import _lib
        from .ops._lib import c_float, c_int, c_ll, ptr, stream_ptr

        lr = self.param_groups[0]["lr"]
        f = self.flat
        if f.params.is_cuda:
            out = f.params if f.params.dtype == _lib.HALF else None
            _lib.call("b2f_sgd_momentum_flat", ptr(self.master), ptr(self.buf), ptr(f.grads),
                      c_int(int(f.grads.dtype == _lib.HALF)), ptr(out), c_ll(f.numel), c_float(lr),
                      c_float(self.momentum), c_float(self.weight_decay), ptr(self.grad_scale), stream_ptr())
            if out is None:
                f.params.copy_(self.master)
        else:
            g = f.grads.float() * (float(self.grad_scale) if self.grad_scale is not None else 1.0)
            self.buf.mul_(self.momentum).add_(g + self.weight_decay * self.master)
            self.master.add_(self.buf, alpha=-lr)
            f.params.copy_(self.master.to(f.params.dtype))

    def zero_grad(self, set_to_none: bool = True):
        self.flat.zero_grad()

    def state_dict(self):
        return {"master": self.master, "momentum_buffer": self.buf, "lr": self.param_groups[0]["lr"]}

    def load_state_dict(self, sd):
        self.master.copy_(sd["master"])
        self.buf.copy_(sd["momentum_buffer"])
        self.flat.params.copy_(self.master.to(self.flat.params.dtype))
