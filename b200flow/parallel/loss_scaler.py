"""Loss scaling for `"fp16": {"enabled": true}` runs.

The reference's Megatron-DeepSpeed examples train in fp16 (`--fp16` and
`"fp16": {"enabled": true}`: examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:33-38,53),
which upstream means DeepSpeed's dynamic loss scaler: the loss is multiplied by a power of two
before backward, gradients are checked for inf/nan on every rank (one MAX all-reduce, SURVEY.md
§2H C6), an overflowing step is skipped and the scale lowered, and the scale is raised again
after a window of clean steps.  The log line carries `loss scale` and the skipped-step count.

Here the state machine is kept with DeepSpeed's defaults and config keys
(`loss_scale`, `initial_scale_power`, `loss_scale_window`, `hysteresis`, `min_loss_scale`,
`consecutive_hysteresis`).  The compute dtype under it stays bf16 — the tcgen05 kernels are
bf16 — so overflows are rare, but a diverging run still trips the same skip logic.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass
class LossScalerConfig:
    static_scale: float = 0.0          # `loss_scale`: 0 = dynamic
    initial_scale_power: int = 16
    scale_window: int = 1000
    hysteresis: int = 2
    min_scale: float = 1.0
    consecutive_hysteresis: bool = False
    scale_factor: float = 2.0

    @classmethod
    def from_ds_config(cls, fp16: dict) -> "LossScalerConfig":
        return cls(static_scale=float(fp16.get("loss_scale", 0) or 0),
                   initial_scale_power=int(fp16.get("initial_scale_power", 16)),
                   scale_window=int(fp16.get("loss_scale_window", 1000)),
                   hysteresis=int(fp16.get("hysteresis", 2)),
                   min_scale=float(fp16.get("min_loss_scale", 1)),
                   consecutive_hysteresis=bool(fp16.get("consecutive_hysteresis", False)))


class LossScaler:
    """Static scale (`"loss_scale": S` with S > 0): overflowing steps are still skipped."""

    dynamic = False

    def __init__(self, scale: float = 1.0):
        self.scale = float(scale)
        self.skipped = 0

    def update(self, overflow: bool) -> None:
        if overflow:
            self.skipped += 1

    def state_dict(self) -> dict:
        return {"cur_scale": self.scale, "skipped": self.skipped, "dynamic": self.dynamic}

    def load_state_dict(self, sd: dict) -> None:
        self.scale = float(sd["cur_scale"])
        self.skipped = int(sd.get("skipped", 0))


class DynamicLossScaler(LossScaler):
    """DeepSpeed `DynamicLossScaler.update_scale` semantics.

    On overflow the scale is divided by `scale_factor` once the hysteresis budget is used up
    (`hysteresis` overflows in a row are tolerated at the start, then every overflow lowers it);
    after `scale_window` steps since the last overflow the scale is multiplied by the factor and
    the hysteresis budget is restored."""

    dynamic = True

    def __init__(self, cfg: LossScalerConfig | None = None):
        cfg = cfg or LossScalerConfig()
        super().__init__(2.0 ** cfg.initial_scale_power)
        self.cfg = cfg
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.cur_hysteresis = cfg.hysteresis

    def update(self, overflow: bool) -> None:
        c = self.cfg
        if overflow:
            self.skipped += 1
            if c.hysteresis == 1 or self.cur_hysteresis == 1:
                self.scale = max(self.scale / c.scale_factor, c.min_scale)
            else:
                self.cur_hysteresis -= 1
            self.last_overflow_iter = self.cur_iter
        else:
            if c.consecutive_hysteresis:
                self.cur_hysteresis = c.hysteresis
            if (self.cur_iter - self.last_overflow_iter) % c.scale_window == 0:
                if not c.consecutive_hysteresis:
                    self.cur_hysteresis = c.hysteresis
                self.scale *= c.scale_factor
        self.cur_iter += 1

    def state_dict(self) -> dict:
        return {**super().state_dict(), "cur_iter": self.cur_iter,
                "last_overflow_iter": self.last_overflow_iter, "cur_hysteresis": self.cur_hysteresis}

    def load_state_dict(self, sd: dict) -> None:
        super().load_state_dict(sd)
        self.cur_iter = int(sd.get("cur_iter", 0))
        self.last_overflow_iter = int(sd.get("last_overflow_iter", -1))
        self.cur_hysteresis = int(sd.get("cur_hysteresis", self.cfg.hysteresis))


def make_loss_scaler(fp16: dict | None):
    """None unless fp16 is enabled in the ds_config section given."""
    if not fp16 or not fp16.get("enabled"):
        return None
    cfg = LossScalerConfig.from_ds_config(fp16)
    return LossScaler(cfg.static_scale) if cfg.static_scale > 0 else DynamicLossScaler(cfg)
