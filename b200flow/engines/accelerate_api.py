"""`Accelerator`-shaped facade over the b200flow runtime.

The reference's BERT workload is Accelerate's `examples/complete_nlp_example.py` launched by
torchrun (examples/accelerate/bert-glue-mrpc/pretrain.yaml:35-51; call stack in SURVEY.md
§3.2).  This class keeps the calls that script makes — `prepare`, `backward`, `accumulate`,
`gather_for_metrics`, `save_state` / `load_state`, `skip_first_batches`, `init_trackers` /
`log` / `end_training`, `main_process_first`, `print`, `wait_for_everyone` — with b200flow
underneath: the model is flattened and wrapped in the bucketed K1 DDP, bf16 parameters with
fp32 master weights in the fused AdamW, checkpoints in Accelerate's directory layout.

`mixed_precision="fp16"` (what the reference passes) is honoured as reduced-precision
compute — bf16 storage on B200, fp32 on CPU — with Accelerate's fp16 behaviour around it: a
`torch.amp.GradScaler`-style dynamic loss scale (65536, x2 after 2000 clean steps, x0.5 on overflow),
optimizer and scheduler steps skipped on overflow (`optimizer_step_was_skipped`), `scaler.pt` in
`save_state`.
"""
from __future__ import annotations

import contextlib
import math
import os

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader

from ..ckpt import accelerate_layout as L
from ..obs.tracking import Tracker
from ..parallel.ddp import DistributedDataParallel


class BatchSamplerShard:
    """Rank r takes batches r, r+N, r+2N, …; the tail is completed by wrapping around so every
    rank sees the same number of batches (Accelerate's `even_batches=True`)."""

    def __init__(self, batch_sampler, num_processes: int, process_index: int):
        self.bs, self.n, self.r = batch_sampler, num_processes, process_index

    def __len__(self):
        return math.ceil(len(self.bs) / self.n)

    def __iter__(self):
        batches = list(self.bs)
        total = len(self) * self.n
        i = 0
        while len(batches) < total:
            batches.append(batches[i % max(1, len(batches))])
            i += 1
        return iter(batches[self.r::self.n])


class _ShardedLoader:
    def __init__(self, dl: DataLoader, acc: "Accelerator"):
        self.acc = acc
        self.total_dataset_length = len(dl.dataset)
        self.batch_size = dl.batch_size
        shard = BatchSamplerShard(dl.batch_sampler, acc.num_processes, acc.process_index)
        self.dl = DataLoader(dl.dataset, batch_sampler=shard, collate_fn=dl.collate_fn,
                             num_workers=0, pin_memory=acc.device.type == "cuda")
        self.skip = 0
        self.end_of_dataloader = False
        self.remainder = self.total_dataset_length % max(1, (dl.batch_size or 1) * acc.num_processes)

    def __len__(self):
        return len(self.dl) - self.skip

    def __iter__(self):
        n = len(self.dl)
        self.end_of_dataloader = False
        for i, batch in enumerate(self.dl):
            if i < self.skip:
                continue
            self.end_of_dataloader = i == n - 1
            self.acc._active_loader = self
            yield {k: (v.to(self.acc.device, non_blocking=True) if isinstance(v, torch.Tensor) else v)
                   for k, v in batch.items()} if isinstance(batch, dict) else batch
        self.skip = 0


class GradScaler:
    """Dynamic loss scale with `torch.amp.GradScaler`'s constants and state-dict keys."""

    def __init__(self, init_scale: float = 65536.0, growth_factor: float = 2.0, backoff_factor: float = 0.5,
                 growth_interval: int = 2000):
        self.scale = float(init_scale)
        self.growth_factor, self.backoff_factor, self.growth_interval = growth_factor, backoff_factor, growth_interval
        self._growth_tracker = 0

    def update(self, found_inf: bool) -> None:
        if found_inf:
            self.scale *= self.backoff_factor
            self._growth_tracker = 0
            return
        self._growth_tracker += 1
        if self._growth_tracker == self.growth_interval:
            self.scale *= self.growth_factor
            self._growth_tracker = 0

    def state_dict(self) -> dict:
        return {"scale": self.scale, "growth_factor": self.growth_factor, "backoff_factor": self.backoff_factor,
                "growth_interval": self.growth_interval, "_growth_tracker": self._growth_tracker}

    def load_state_dict(self, sd: dict) -> None:
        self.scale = float(sd["scale"])
        self.growth_factor, self.backoff_factor = float(sd["growth_factor"]), float(sd["backoff_factor"])
        self.growth_interval, self._growth_tracker = int(sd["growth_interval"]), int(sd["_growth_tracker"])


class _Optimizer:
    """AcceleratedOptimizer: defers to the wrapped optimizer, finishing the DDP reduction first
    and honouring gradient accumulation; under fp16 it unscales, checks for inf / nan and skips."""

    def __init__(self, opt, acc):
        self.optimizer, self.acc = opt, acc
        self.step_was_skipped = False
        # the DDP hooks move every gradient into `p.main_grad` (fp32, flat) and clear `p.grad`; b200flow's own
        # optimizers read `main_grad`, a stock torch.optim optimizer (what the Accelerate example constructs:
        # AdamW) reads `p.grad` and would silently step on nothing
        self._reads_main_grad = hasattr(opt, "_grad") or hasattr(opt, "flat")

    def _expose_grads(self):
        """Hand the reduced `main_grad` to an optimizer that only knows `p.grad`."""
        for g in self.optimizer.param_groups:
            for p in g["params"]:
                mg = getattr(p, "main_grad", None)
                if mg is not None and p.requires_grad:
                    p.grad = mg.to(p.dtype) if mg.dtype != p.dtype else mg

    def _step_inner(self):
        if not self._reads_main_grad:
            self._expose_grads()
        self.optimizer.step()
        if not self._reads_main_grad:
            for g in self.optimizer.param_groups:
                for p in g["params"]:
                    if getattr(p, "main_grad", None) is not None:
                        p.grad = None

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    def _grads(self):
        seen, out = set(), []
        for g in self.optimizer.param_groups:
            for p in g["params"]:
                owner = getattr(p, "_flat_owner", None)
                if owner is not None:
                    if id(owner) not in seen:
                        seen.add(id(owner))
                        out.append(owner.grads)
                    continue
                t = getattr(p, "main_grad", None)
                t = t if t is not None else p.grad
                if t is not None:
                    out.append(t)
        return out

    def step(self):
        if not self.acc.sync_gradients:
            return
        scaler = self.acc.scaler
        if scaler is None:
            self._step_inner()
            return
        grads = self._grads()
        found_inf = not all(bool(torch.isfinite(g).all()) for g in grads)    # reduced gradients: same on every rank
        self.step_was_skipped = found_inf
        if not found_inf:
            inv = 1.0 if self.acc._unscaled else 1.0 / scaler.scale
            if hasattr(self.optimizer, "grad_scale"):      # fused AdamW: the unscale rides on its gradient scale
                prev = self.optimizer.grad_scale
                self.optimizer.grad_scale = torch.full((1,), inv, dtype=torch.float32, device=grads[0].device) \
                    if prev is None else prev * inv
                self.optimizer.step()
                self.optimizer.grad_scale = prev
            else:
                for g in grads:
                    g.mul_(inv)
                self._step_inner()
        self.acc._unscaled = False
        scaler.update(found_inf)

    def zero_grad(self, set_to_none=True):
        if self.acc.sync_gradients:
            self.optimizer.zero_grad()
            if not self._reads_main_grad:      # a stock optimizer clears p.grad only; the flat buffer is ours
                for m in self.acc._models:
                    m.zero_grad()

    def state_dict(self):
        return self.optimizer.state_dict()

    def load_state_dict(self, sd):
        self.optimizer.load_state_dict(sd)


class _Scheduler:
    def __init__(self, sch, acc):
        self.scheduler, self.acc = sch, acc

    def step(self):
        if self.acc.sync_gradients and not self.acc.optimizer_step_was_skipped:
            self.scheduler.step()

    def get_last_lr(self):
        return self.scheduler.get_last_lr()

    def state_dict(self):
        return self.scheduler.state_dict()

    def load_state_dict(self, sd):
        self.scheduler.load_state_dict(sd)


class Accelerator:
    def __init__(self, cpu: bool = False, mixed_precision: str | None = None, log_with=None,
                 project_dir: str | None = None, gradient_accumulation_steps: int = 1,
                 bucket_cap_mb: float = 25.0, fused_collectives: bool = True):
        self.mixed_precision = mixed_precision or "no"
        self.project_dir = project_dir
        self.log_with = log_with
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.bucket_cap_mb, self.fused = bucket_cap_mb, fused_collectives
        use_cuda = torch.cuda.is_available() and not cpu
        if "RANK" in os.environ and not dist.is_initialized():
            from ..parallel.groups import init_distributed

            init_distributed("nccl" if use_cuda else "gloo")
        self.num_processes = dist.get_world_size() if dist.is_initialized() else 1
        self.process_index = dist.get_rank() if dist.is_initialized() else 0
        self.local_process_index = int(os.environ.get("LOCAL_RANK", 0))
        self.device = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
        self.param_dtype = torch.bfloat16 if use_cuda else torch.float32
        if use_cuda:       # mixed_precision="fp16" (the reference's --mixed_precision fp16): the fp16 build of the kernels
            from ..ops import _lib

            _lib.set_half("fp16" if self.mixed_precision == "fp16" else "bf16")
            self.param_dtype = _lib.HALF
        self.scaler = GradScaler() if self.mixed_precision == "fp16" else None
        self._unscaled = False
        self.trackers: list[Tracker] = []
        self._models, self._optimizers, self._schedulers, self._loaders = [], [], [], []
        self._custom = []
        self._accum = 0
        self.sync_gradients = True
        self.step = 0
        self._active_loader = None

    # ----------------------------------------------------------------- props
    @property
    def is_main_process(self) -> bool:
        return self.process_index == 0

    @property
    def is_local_main_process(self) -> bool:
        return self.local_process_index == 0

    @property
    def use_distributed(self) -> bool:
        return self.num_processes > 1

    def print(self, *a, **kw):
        if self.is_main_process:
            print(*a, **kw, flush=True)

    def wait_for_everyone(self):
        if self.use_distributed:
            dist.barrier()

    @contextlib.contextmanager
    def main_process_first(self):
        if not self.is_main_process:
            self.wait_for_everyone()
        yield
        if self.is_main_process:
            self.wait_for_everyone()

    # --------------------------------------------------------------- prepare
    def prepare(self, *objs):
        out = []
        for o in objs:
            if isinstance(o, torch.nn.Module):
                o = o.to(self.device)
                if any(p.is_floating_point() and p.dtype != self.param_dtype for p in o.parameters()):
                    o = o.to(self.param_dtype)
                ddp = DistributedDataParallel(o, bucket_cap_mb=self.bucket_cap_mb, fused=self.fused) \
                    if True else o
                self._models.append(ddp)
                out.append(ddp)
            elif isinstance(o, torch.optim.Optimizer):
                w = _Optimizer(o, self)
                self._optimizers.append(w)
                out.append(w)
            elif isinstance(o, DataLoader):
                w = _ShardedLoader(o, self)
                self._loaders.append(w)
                out.append(w)
            elif hasattr(o, "step") and hasattr(o, "state_dict"):
                w = _Scheduler(o, self)
                self._schedulers.append(w)
                out.append(w)
            else:
                out.append(o)
        return tuple(out) if len(out) != 1 else out[0]

    def unwrap_model(self, model):
        return model.module if isinstance(model, DistributedDataParallel) else model

    def register_for_checkpointing(self, *objs):
        self._custom.extend(objs)

    # -------------------------------------------------------------- training
    @contextlib.contextmanager
    def accumulate(self, *models):
        self._accum += 1
        self.sync_gradients = (self._accum % self.gradient_accumulation_steps == 0) or (
            self._active_loader is not None and self._active_loader.end_of_dataloader)
        ctxs = [m.no_sync() for m in self._models if not self.sync_gradients]
        with contextlib.ExitStack() as st:
            for c in ctxs:
                st.enter_context(c)
            yield

    @property
    def optimizer_step_was_skipped(self) -> bool:
        return any(o.step_was_skipped for o in self._optimizers)

    def backward(self, loss, **kw):
        if self.gradient_accumulation_steps > 1:
            loss = loss / self.gradient_accumulation_steps
        if self.scaler is not None:
            loss = loss * self.scaler.scale
        loss.backward(**kw)
        if self.sync_gradients:
            for m in self._models:
                m.finish_reduction()
            self.step += 1

    def unscale_gradients(self) -> None:
        """Divide the (reduced) gradients by the loss scale once per step (fp16 only)."""
        if self.scaler is None or self._unscaled:
            return
        for o in self._optimizers:
            for g in o._grads():
                g.mul_(1.0 / self.scaler.scale)
        self._unscaled = True

    def clip_grad_norm_(self, parameters, max_norm: float, norm_type: float = 2.0):
        self.unscale_gradients()
        grads = [getattr(p, "main_grad", None) if getattr(p, "main_grad", None) is not None else p.grad
                 for p in parameters]
        grads = [g for g in grads if g is not None]
        total = torch.sqrt(sum(g.float().pow(2).sum() for g in grads))
        coef = (max_norm / (total + 1e-6)).clamp(max=1.0)
        for g in grads:
            g.mul_(coef.to(g.dtype))
        return total

    # --------------------------------------------------------------- metrics
    def gather(self, t: torch.Tensor) -> torch.Tensor:
        if not self.use_distributed:
            return t
        t = t.contiguous()
        # ranks may hold different batch sizes on the last step: pad to the max, trim after
        n = torch.tensor([t.shape[0]], device=t.device)
        sizes = [torch.zeros_like(n) for _ in range(self.num_processes)]
        dist.all_gather(sizes, n)
        m = int(max(s.item() for s in sizes))
        if t.shape[0] < m:
            t = torch.cat([t, t.new_zeros((m - t.shape[0], *t.shape[1:]))], 0)
        out = [torch.empty_like(t) for _ in range(self.num_processes)]
        dist.all_gather(out, t)
        # rank r held batches r, r+N, ...: concatenating by rank restores dataset order
        return torch.cat([o[:int(s.item())] for o, s in zip(out, sizes)], 0)

    def gather_for_metrics(self, tensors):
        """all_gather predictions/labels and drop the samples duplicated to even out the last
        batch."""
        single = isinstance(tensors, torch.Tensor)
        ts = [tensors] if single else list(tensors)
        out = [self.gather(t) for t in ts]
        ld = self._active_loader
        if ld is not None and ld.end_of_dataloader and ld.remainder > 0 and self.use_distributed:
            out = [o[:ld.remainder] if o.shape[0] > ld.remainder else o for o in out]
        return out[0] if single else tuple(out)

    def reduce(self, t: torch.Tensor, reduction: str = "sum", scale: float = 1.0) -> torch.Tensor:
        """`accelerator.reduce`: all-reduce a tensor over the processes ("sum" or "mean"), times `scale`."""
        out = t.clone()
        if self.use_distributed:
            dist.all_reduce(out)
            if reduction == "mean":
                out = out / self.num_processes
        return out * scale if scale != 1.0 else out

    def pad_across_processes(self, t: torch.Tensor, dim: int = 0, pad_index: int = 0, pad_first: bool = False):
        """Pad `t` along `dim` to the largest size any process holds, so that it can be gathered."""
        if not self.use_distributed:
            return t
        n = torch.tensor([t.shape[dim]], device=t.device)
        sizes = [torch.zeros_like(n) for _ in range(self.num_processes)]
        dist.all_gather(sizes, n)
        m = int(max(s.item() for s in sizes))
        if m == t.shape[dim]:
            return t
        shape = list(t.shape)
        shape[dim] = m - t.shape[dim]
        pad = t.new_full(shape, pad_index)
        return torch.cat([pad, t] if pad_first else [t, pad], dim)

    @contextlib.contextmanager
    def autocast(self):
        """Mixed-precision region.  Parameters already live in the compute dtype (bf16 on B200), so this only
        wraps CPU-side or stray fp32 modules in torch's autocast."""
        if self.device.type == "cuda" and self.mixed_precision in ("fp16", "bf16"):
            with torch.autocast("cuda", dtype=self.param_dtype):
                yield
        else:
            yield

    def on_main_process(self, fn):
        """Decorator: run `fn` on the main process only."""
        def wrapped(*a, **kw):
            return fn(*a, **kw) if self.is_main_process else None
        return wrapped

    @property
    def is_last_process(self) -> bool:
        return self.process_index == self.num_processes - 1

    def save_model(self, model, save_directory: str, safe_serialization: bool = True) -> None:
        """Weights of one model in Accelerate's file naming (`model.safetensors` / `pytorch_model.bin`)."""
        self.wait_for_everyone()
        if self.is_main_process:
            os.makedirs(save_directory, exist_ok=True)
            L.save_model_weights(self.unwrap_model(model), save_directory, safe_serialization)
        self.wait_for_everyone()

    def get_tracker(self, name: str = "all"):
        return self.trackers[0] if self.trackers else None

    # -------------------------------------------------------------- tracking
    def init_trackers(self, project_name: str, config: dict | None = None):
        if self.is_main_process and self.log_with is not None:
            self.trackers.append(Tracker(self.project_dir or ".", project_name, config))

    def log(self, values: dict, step: int | None = None):
        for t in self.trackers:
            t.log(values, step)

    def end_training(self):
        for t in self.trackers:
            t.finish()
        self.wait_for_everyone()

    # ----------------------------------------------------------- checkpoints
    def save_state(self, output_dir: str, safe_serialization: bool = True) -> str:
        self.wait_for_everyone()
        models = [self.unwrap_model(m) for m in self._models]
        L.save_state(output_dir, models, [o.optimizer for o in self._optimizers],
                     [s.scheduler for s in self._schedulers], [], self.scaler, self.process_index,
                     self.is_main_process, self._custom, safe_serialization)
        self.wait_for_everyone()
        return output_dir

    def load_state(self, input_dir: str) -> None:
        models = [self.unwrap_model(m) for m in self._models]
        L.load_state(input_dir, models, [o.optimizer for o in self._optimizers],
                     [s.scheduler for s in self._schedulers], self.scaler, self.process_index, self._custom)

    def skip_first_batches(self, dataloader, num_batches: int):
        dataloader.skip = num_batches
        return dataloader

    def free_memory(self):
        self._models, self._optimizers, self._schedulers, self._loaders = [], [], [], []
