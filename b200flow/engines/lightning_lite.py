"""The slice of PyTorch-Lightning the reference's Ray workload uses (`LightningModule`,
`Trainer.fit`, callbacks, `self.log`), on the b200flow DDP (RayDDPStrategy's role) — see
examples/ray/lightning-bert/fine-tune.yaml:17 (pytorch-lightning==2.2.1 pinned there)."""
from __future__ import annotations

import os
import tempfile

import torch
import torch.distributed as dist
from torch import nn

from ..parallel.ddp import DistributedDataParallel


class LightningModule(nn.Module):
    def __init__(self):
        super().__init__()
        self._logged: dict = {}
        self.trainer = None

    def log(self, name, value, sync_dist: bool = False, **kw):
        v = float(value.detach()) if isinstance(value, torch.Tensor) else float(value)
        if sync_dist and dist.is_initialized() and dist.get_world_size() > 1:
            dev = next(self.parameters()).device
            t = torch.tensor([v], device=dev if dist.get_backend() == "nccl" else "cpu")
            dist.all_reduce(t)
            v = float(t) / dist.get_world_size()
        self._logged[name] = v


class Callback:
    def on_train_epoch_end(self, trainer, module):
        pass


class RayTrainReportCallback(Callback):
    """Report the logged metrics and a checkpoint of the module at the end of every epoch."""

    def on_train_epoch_end(self, trainer, module):
        from . import ray_train_api as train

        metrics = dict(module._logged, epoch=trainer.current_epoch, step=trainer.global_step)
        tmp = tempfile.mkdtemp(prefix="b2f_ckpt_")
        if train.get_context().get_world_rank() == 0:
            torch.save({"state_dict": trainer.unwrapped.state_dict(), "epoch": trainer.current_epoch},
                       os.path.join(tmp, "checkpoint.ckpt"))
        train.report(metrics, checkpoint=train.Checkpoint.from_directory(tmp))


def _unpack_optimizers(cfg):
    """Lightning's `configure_optimizers` return shapes: an optimizer; a list of them; (optimizers, schedulers);
    {"optimizer": ..., "lr_scheduler": scheduler | {"scheduler": ..., "interval": "epoch" | "step"}}.
    Returns (optimizer, scheduler | None, interval)."""
    sched, interval = None, "epoch"
    if isinstance(cfg, dict):
        opt = cfg["optimizer"]
        ls = cfg.get("lr_scheduler")
        if isinstance(ls, dict):
            sched, interval = ls["scheduler"], ls.get("interval", "epoch")
        else:
            sched = ls
    elif isinstance(cfg, tuple) and len(cfg) == 2 and isinstance(cfg[0], (list, tuple)):
        opt = cfg[0][0]
        first = cfg[1][0] if cfg[1] else None
        if isinstance(first, dict):
            sched, interval = first["scheduler"], first.get("interval", "epoch")
        else:
            sched = first
    elif isinstance(cfg, (list, tuple)):
        opt = cfg[0]
    else:
        opt = cfg
    return opt, sched, interval


class Trainer:
    """`pl.Trainer` for the calls the reference's workload makes, plus the common loop controls: `max_epochs`,
    `max_steps`, `accumulate_grad_batches`, `gradient_clip_val`, `limit_train_batches`, `limit_val_batches`."""

    def __init__(self, max_epochs: int = 1, accelerator: str = "auto", devices="auto", strategy=None,
                 plugins=None, callbacks=None, enable_progress_bar: bool = False, max_steps: int = -1,
                 accumulate_grad_batches: int = 1, gradient_clip_val: float | None = None,
                 limit_train_batches: int | None = None, limit_val_batches: int | None = None, **kw):
        self.max_epochs, self.callbacks = max_epochs, list(callbacks or [])
        self.max_steps, self.accumulate = max_steps, max(1, int(accumulate_grad_batches))
        self.clip, self.limit_train, self.limit_val = gradient_clip_val, limit_train_batches, limit_val_batches
        self.current_epoch = 0
        self.global_step = 0
        self.unwrapped = None

    def _clip(self, model):
        grads = [getattr(p, "main_grad", None) if getattr(p, "main_grad", None) is not None else p.grad
                 for p in model.parameters()]
        grads = [g for g in grads if g is not None]
        if not grads:
            return
        total = torch.sqrt(sum(g.float().pow(2).sum() for g in grads))
        coef = (self.clip / (total + 1e-6)).clamp(max=1.0)
        for g in grads:
            g.mul_(coef.to(g.dtype))

    def fit(self, model: LightningModule, train_dataloaders=None, val_dataloaders=None):
        use_cuda = torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"
        dev = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
        model.to(dev)
        if use_cuda:
            model.to(torch.bfloat16)
        model.trainer = self
        self.unwrapped = model
        ddp = DistributedDataParallel(model) if dist.is_initialized() else None
        opt, sched, interval = _unpack_optimizers(model.configure_optimizers())
        to_dev = lambda b: {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in b.items()}  # noqa: E731
        done = False
        for epoch in range(self.max_epochs):
            self.current_epoch = epoch
            model.train()
            pending = 0
            for i, batch in enumerate(train_dataloaders() if callable(train_dataloaders) else train_dataloaders):
                if self.limit_train is not None and i >= self.limit_train:
                    break
                last_micro = (i + 1) % self.accumulate == 0
                loss = model.training_step(to_dev(batch), i)
                ctx = ddp.no_sync() if (ddp is not None and not last_micro) else None
                if ctx is not None:
                    with ctx:
                        (loss / self.accumulate).backward()
                else:
                    (loss / self.accumulate).backward()
                pending += 1
                if not last_micro:
                    continue
                if ddp is not None:
                    ddp.finish_reduction()
                if self.clip:
                    self._clip(model)
                opt.step()
                opt.zero_grad()
                pending = 0
                self.global_step += 1
                if sched is not None and interval == "step":
                    sched.step()
                if 0 < self.max_steps <= self.global_step:
                    done = True
                    break
            if pending and not done:        # a trailing partial accumulation window still updates (Lightning does)
                if ddp is not None:
                    ddp.require_sync = True
                    ddp.finish_reduction()
                if self.clip:
                    self._clip(model)
                opt.step()
                opt.zero_grad()
                self.global_step += 1
            if sched is not None and interval == "epoch":
                sched.step()
            if val_dataloaders is not None:
                model.eval()
                with torch.no_grad():
                    for i, batch in enumerate(val_dataloaders() if callable(val_dataloaders) else val_dataloaders):
                        if self.limit_val is not None and i >= self.limit_val:
                            break
                        model.validation_step(to_dev(batch), i)
                if hasattr(model, "on_validation_epoch_end"):
                    model.on_validation_epoch_end()
            if hasattr(model, "on_train_epoch_end"):
                model.on_train_epoch_end()
            for cb in self.callbacks:
                cb.on_train_epoch_end(self, model)
            if done:
                break


class RayDDPStrategy:
    """`ray.train.lightning.RayDDPStrategy`: in the reference's script it tells Lightning to use the process
    group Ray Train already created.  Here `Trainer.fit` always wraps the module in b200flow's DDP on the
    existing process group, so the object only carries its options (`find_unused_parameters` is not needed:
    the reducer sweeps buckets whose parameters received no gradient)."""

    def __init__(self, find_unused_parameters: bool = False, **kw):
        self.find_unused_parameters, self.options = find_unused_parameters, kw


class RayLightningEnvironment:
    """`ray.train.lightning.RayLightningEnvironment`: rank / world size as the worker group set them."""

    def world_size(self) -> int:
        return dist.get_world_size() if dist.is_initialized() else 1

    def global_rank(self) -> int:
        return dist.get_rank() if dist.is_initialized() else 0

    def local_rank(self) -> int:
        return int(os.environ.get("LOCAL_RANK", 0))

    def node_rank(self) -> int:
        return 0


def prepare_trainer(trainer: Trainer) -> Trainer:
    """`ray.train.lightning.prepare_trainer`: validates that the trainer is set up for the worker group; the
    local trainer has nothing to patch, so it is returned as it is."""
    return trainer
