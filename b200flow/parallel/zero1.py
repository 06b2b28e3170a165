"""ZeRO stage-1 optimizer: optimizer state partitioned over the data-parallel group.

What DeepSpeed's `DeepSpeedZeroOptimizer` (stage 1) does for the reference's
`"zero_optimization": {"stage": 1}` (examples/megatron-deepspeed/gpt2_345m/
pretrain-ddp-zero1.yaml:33-38), re-designed around one
flat buffer per model chunk:

    grads (fp32 flat)  --reduce-scatter-->  owner's shard
    shard: unscale / clip -> AdamW on fp32 master -> bf16 working copy
    params (bf16 flat) <--all-gather--      every owner's shard

Two execution paths share this class: the *fused* K2 kernel (`parallel/fused_dp.py`,
one launch doing reduce-scatter + Adam + all-gather over NVLink peer memory) and
the *baseline* path (NCCL/gloo reduce-scatter, fused-Adam kernel, all-gather).
On CPU the Adam math falls back to the oracle so the partition logic is testable
under gloo.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.distributed as dist

from ..ops import optim as O
from ..ops import ref
from ..utils.flat import FlatParams, partition_range


@dataclass
class AdamConfig:
    lr: float = 1.5e-4
    beta1: float = 0.9
    beta2: float = 0.95
    eps: float = 1e-8
    weight_decay: float = 0.01
    clip_grad: float = 1.0


class Zero1Optimizer:
    def __init__(self, flat: FlatParams, cfg: AdamConfig, dp_group=None, dp_world: int = 1,
                 dp_rank: int = 0, norm_groups: list | None = None, fused=None,
                 grad_divisor: float | None = None):
        assert flat.numel % dp_world == 0, "pad the flat buffer to a multiple of the DP size"
        self.flat, self.cfg = flat, cfg
        self.dp_group, self.dp_world, self.dp_rank = dp_group, dp_world, dp_rank
        self.norm_groups = norm_groups or []   # extra groups (tp, pp) to sum the grad norm over
        self.fused = fused
        self.lo, self.hi = partition_range(flat.numel, dp_world, dp_rank)
        n = self.hi - self.lo
        dev = flat.params.device
        self.master = flat.params[self.lo:self.hi].float().clone()
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.grad_shard = torch.zeros(n, dtype=torch.float32, device=dev) if dp_world > 1 else None
        self.step_count = 0
        self.lr = cfg.lr
        self.grad_divisor = grad_divisor  # micro-batches x dp (mean over the global batch)
        self.last_grad_norm = None
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=dev)

    # ------------------------------------------------------------------ step
    def step(self, grad_divisor: float | None = None) -> None:
        cfg, flat = self.cfg, self.flat
        div = grad_divisor if grad_divisor is not None else (self.grad_divisor or 1.0)
        self.step_count += 1
        if self.fused is not None and flat.grads.is_cuda and self.dp_world > 1:
            self.last_grad_norm = self.fused.step(self, div)
            return
        # 1. reduce-scatter (sum) the fp32 grads over DP
        if self.dp_world > 1:
            if flat.grads.is_cuda:
                dist.reduce_scatter_tensor(self.grad_shard, flat.grads, group=self.dp_group)
            else:
                dist.all_reduce(flat.grads, group=self.dp_group)
                self.grad_shard.copy_(flat.grads[self.lo:self.hi])
            g = self.grad_shard
        else:
            g = flat.grads
        # 2. global grad norm (of the averaged gradient) -> clip coefficient on device
        if g.is_cuda:
            O.sumsq(g, out=self._sumsq)
        else:
            self._sumsq[0] = g.double().pow(2).sum().float()
        if self.dp_world > 1:
            dist.all_reduce(self._sumsq, group=self.dp_group)
        for grp in self.norm_groups:
            if grp is not None:
                dist.all_reduce(self._sumsq, group=grp)
        inv = 1.0 / div
        if g.is_cuda:
            # norm of averaged grad = inv * sqrt(sumsq); fold `inv` into the scales
            self._sumsq.mul_(inv * inv)
            coef, nrm = O.clip_coef(self._sumsq, cfg.clip_grad, extra_scale=inv)
            self.last_grad_norm = nrm
            O.adamw_flat_(self.master, self.exp_avg, self.exp_avg_sq, g,
                          flat.params[self.lo:self.hi], lr=self.lr, beta1=cfg.beta1,
                          beta2=cfg.beta2, eps=cfg.eps, weight_decay=cfg.weight_decay,
                          step=self.step_count, grad_scale=coef)
        else:
            nrm = math.sqrt(float(self._sumsq[0])) * inv
            coef = min(1.0, cfg.clip_grad / (nrm + 1e-6)) if cfg.clip_grad > 0 else 1.0
            self.last_grad_norm = torch.tensor([nrm])
            ref.adamw_step(self.master, self.exp_avg, self.exp_avg_sq, g * (coef * inv), lr=self.lr,
                           beta1=cfg.beta1, beta2=cfg.beta2, eps=cfg.eps,
                           weight_decay=cfg.weight_decay, step=self.step_count)
            flat.params[self.lo:self.hi].copy_(self.master.to(flat.params.dtype))
        # 3. all-gather the updated bf16 shards
        if self.dp_world > 1:
            dist.all_gather_into_tensor(flat.params, flat.params[self.lo:self.hi].clone(),
                                        group=self.dp_group)

    def zero_grad(self) -> None:
        self.flat.zero_grad()

    # ------------------------------------------------------------ checkpoint
    def state_dict(self) -> dict:
        return {"step": self.step_count, "lr": self.lr, "partition": (self.lo, self.hi),
                "dp_world": self.dp_world, "dp_rank": self.dp_rank,
                "fp32_flat": self.master, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq}

    def load_state_dict(self, sd: dict) -> None:
        assert tuple(sd["partition"]) == (self.lo, self.hi), "checkpoint was written with another DP layout"
        self.step_count, self.lr = sd["step"], sd["lr"]
        self.master.copy_(sd["fp32_flat"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.flat.params[self.lo:self.hi].copy_(self.master.to(self.flat.params.dtype))


def cosine_lr(step: int, *, base_lr: float, min_lr: float, warmup_steps: int, decay_steps: int) -> float:
    """Megatron's `--lr-decay-style cosine` with linear warm-up
    (examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:47-52)."""
    if warmup_steps > 0 and step <= warmup_steps:
        return base_lr * step / warmup_steps
    if step >= decay_steps:
        return min_lr
    frac = (step - warmup_steps) / max(1, decay_steps - warmup_steps)
    return min_lr + 0.5 * (base_lr - min_lr) * (1.0 + math.cos(math.pi * frac))


def megatron_lr(step: int, *, style: str = "cosine", base_lr: float, min_lr: float, warmup_steps: int,
                decay_steps: int) -> float:
    """Megatron's `OptimizerParamScheduler.get_lr` for `--lr-decay-style` constant / linear / cosine /
    inverse-square-root: linear warm-up to `base_lr`, then the decay between `base_lr` and `min_lr` over
    `decay_steps`, `min_lr` afterwards."""
    if warmup_steps > 0 and step <= warmup_steps:
        return base_lr * step / warmup_steps
    if style == "constant":
        return base_lr
    if style == "inverse-square-root":
        return max(min_lr, base_lr * math.sqrt(max(warmup_steps, 1)) / math.sqrt(max(step, 1)))
    if step >= decay_steps:
        return min_lr
    frac = (step - warmup_steps) / max(1, decay_steps - warmup_steps)
    if style == "linear":
        coeff = 1.0 - frac
    elif style == "cosine":
        coeff = 0.5 * (1.0 + math.cos(math.pi * frac))
    else:
        raise ValueError(f"unknown --lr-decay-style {style!r}")
    return min_lr + coeff * (base_lr - min_lr)
