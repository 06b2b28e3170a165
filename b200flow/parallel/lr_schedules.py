"""DeepSpeed's learning-rate schedules, selected by the `"scheduler": {"type": ..., "params": {...}}` block of a
ds_config (the reference writes its ds_config inline and lets Megatron drive the rate —
examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:33-38 — but the engine's users may name one):

    WarmupLR        warmup_min_lr -> warmup_max_lr over warmup_num_steps ("log" or "linear"), then constant
    WarmupDecayLR   the same warm-up, then linear decay to 0 at total_num_steps
    WarmupCosineLR  linear warm-up from warmup_min_ratio x lr, then cosine to cos_min_ratio x lr at total_num_steps
    (no type)       Megatron's schedule (`parallel/zero1.py::megatron_lr`): max_lr, min_lr, warmup_steps, decay_steps

Every schedule is a pure function step -> rate (step counts optimizer steps from 1), so a resumed run needs no
scheduler state beyond the step count.
"""
from __future__ import annotations

import math

from .zero1 import megatron_lr


def _warmup_factor(step: int, n: int, kind: str) -> float:
    if n <= 0 or step >= n:
        return 1.0
    if kind == "log":
        return math.log(step + 1) / math.log(n + 1) if n > 0 else 1.0
    return step / n


def make_schedule(block: dict | None, base_lr: float):
    """Returns f(step) -> lr, or None when the config names no schedule."""
    if not block:
        return None
    kind, p = block.get("type"), dict(block.get("params") or {})
    if kind in (None, "Megatron"):
        if "max_lr" not in p:
            return None
        return lambda step: megatron_lr(step, style=str(p.get("decay_style", "cosine")), base_lr=float(p["max_lr"]),
                                        min_lr=float(p.get("min_lr", 0.0)), warmup_steps=int(p.get("warmup_steps", 0)),
                                        decay_steps=int(p.get("decay_steps", 1)))
    if kind in ("WarmupLR", "WarmupDecayLR"):
        lo, hi = float(p.get("warmup_min_lr", 0.0)), float(p.get("warmup_max_lr", base_lr))
        n, wtype = int(p.get("warmup_num_steps", 1000)), str(p.get("warmup_type", "log"))
        total = int(p.get("total_num_steps", 0))
        if kind == "WarmupDecayLR":
            assert total > n, "WarmupDecayLR: total_num_steps must exceed warmup_num_steps"

        def f(step):
            if step < n:
                return lo + (hi - lo) * _warmup_factor(step, n, wtype)
            if kind == "WarmupLR":
                return hi
            return hi * max(0.0, (total - step) / max(1, total - n))
        return f
    if kind == "WarmupCosineLR":
        total, n = int(p["total_num_steps"]), int(p.get("warmup_num_steps", 1000))
        wmin, cmin = float(p.get("warmup_min_ratio", 0.0)), float(p.get("cos_min_ratio", 1e-4))

        def g(step):
            if step < n:
                return base_lr * (wmin + (1.0 - wmin) * step / max(1, n))
            frac = min(1.0, (step - n) / max(1, total - n))
            return base_lr * (cmin + (1.0 - cmin) * 0.5 * (1.0 + math.cos(math.pi * frac)))
        return g
    raise ValueError(f"ds_config scheduler type {kind!r} is not known (WarmupLR, WarmupDecayLR, WarmupCosineLR)")
