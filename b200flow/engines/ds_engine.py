"""DeepSpeed-shaped training engine over the b200flow runtime.

API surface kept from DeepSpeed so Megatron-DeepSpeed style scripts port 1:1
(the reference drives it as `deepspeed.initialize(...)` + `engine.train_batch()`
from `pretrain_gpt.py --deepspeed --deepspeed_config ds_config.json`,
examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml:33-38,74-85):

    engine = initialize(model_config=cfg, config=ds_config, parallel=ParallelState|None)
    loss = engine.train_batch(data_iter)      # GAS micro-batches + optimizer step
    engine.save_checkpoint(dir) / engine.load_checkpoint(dir)

Recognised ds_config keys: train_micro_batch_size_per_gpu, gradient_accumulation_steps,
zero_optimization.stage (0/1), bf16.enabled, fp16.{enabled, loss_scale, initial_scale_power,
loss_scale_window, hysteresis, min_loss_scale} (dynamic loss scaling with skipped steps; the
compute dtype stays bf16), gradient_clipping,
optimizer.params.{lr,betas,eps,weight_decay}, wall_clock_breakdown.  Unknown keys warn.

Internals are b200flow's own: flat bf16 params + fp32 main_grad, tcgen05 kernels,
ZeRO-1 over the DP group, TP+SP through TPContext, 1F1B over NCCL p2p.
"""
from __future__ import annotations

import os
import warnings
from typing import Iterator

import torch
import torch.distributed as dist

from ..models.gpt import GPTConfig, GPTModel
from ..parallel import pipeline as PP
from ..parallel.groups import ParallelState
from ..parallel.loss_scaler import make_loss_scaler
from ..parallel.tensor_parallel import SINGLE, TPContext, allreduce_replicated_grads
from ..parallel.zero1 import AdamConfig, Zero1Optimizer
from ..utils.flat import FlatParams

_KNOWN = {"train_micro_batch_size_per_gpu", "gradient_accumulation_steps", "zero_optimization",
          "bf16", "fp16", "gradient_clipping", "optimizer", "wall_clock_breakdown",
          "train_batch_size", "steps_per_print", "scheduler", "fused_collectives",
          "activation_checkpointing", "sequence_parallel", "expert_parallel_size", "flops_profiler",
          "comms_logger", "sequence_parallel_size", "checkpoint", "pipeline"}


def split_layers(num_layers: int, stages: int) -> list[tuple[int, int]]:
    """Uniform layer partition (DeepSpeed `partition_method="uniform"`)."""
    base, rem = divmod(num_layers, stages)
    out, lo = [], 0
    for s in range(stages):
        hi = lo + base + (1 if s < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


class TrainEngine:
    def __init__(self, model_config: GPTConfig, config: dict, parallel: ParallelState | None = None,
                 device=None, seed: int = 1234):
        for k in config:
            if k not in _KNOWN:
                warnings.warn(f"ds_config key {k!r} is not interpreted by b200flow (ignored)")
        self.cfg = model_config
        self.ds = config
        self.ps = parallel
        self.device = device if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available()
            else torch.device("cpu"))
        self.micro_batch = int(config.get("train_micro_batch_size_per_gpu", 1))
        self.gas = int(config.get("gradient_accumulation_steps", 1))
        zero_stage = int(config.get("zero_optimization", {}).get("stage", 1))
        self._train_batch_size = config.get("train_batch_size")      # checked against micro x accumulation x DP below
        assert zero_stage in (0, 1), "b200flow implements ZeRO stage 0/1 (what the reference uses)"
        dtype = torch.bfloat16 if self.device.type == "cuda" else torch.float32
        if config.get("bf16", {}).get("enabled") and self.device.type == "cpu":
            dtype = torch.bfloat16
        if self.device.type == "cuda":
            # "fp16": {"enabled": true} (both GPT-2 values files of the reference, with --fp16): IEEE fp16 storage and
            # tensor-core operands, loss scaling below; anything else: bf16.  The choice is process-wide (ops/_lib.py).
            from ..ops import _lib

            want = "fp16" if (config.get("fp16") or {}).get("enabled") and not (config.get("bf16") or {}).get("enabled") \
                else "bf16"
            if os.environ.get("B2F_FP16_AS_BF16", "0") == "1":      # round-1 behaviour: keep the loss scaler, compute in bf16
                want = "bf16"
            _lib.set_half(want)
            dtype = _lib.HALF
        self.dtype = dtype
        use_fused = bool(config.get("fused_collectives", True)) and self.device.type == "cuda"

        ps = parallel
        tp_world = ps.topo.tp if ps else 1
        pp_world = ps.topo.pp if ps else 1
        dp_world = ps.topo.dp if ps else 1
        self.tp_world, self.pp_world, self.dp_world = tp_world, pp_world, dp_world
        if self._train_batch_size is not None:
            # DeepSpeed's rule: train_batch_size = micro batch x accumulation steps x data-parallel size; any one of
            # the three may be left out of the config and is derived from the other two
            tbs = int(self._train_batch_size)
            if "gradient_accumulation_steps" not in config and "train_micro_batch_size_per_gpu" in config:
                self.gas = max(1, tbs // (self.micro_batch * dp_world))
            elif "train_micro_batch_size_per_gpu" not in config:
                self.micro_batch = max(1, tbs // (self.gas * dp_world))
            if self.micro_batch * self.gas * dp_world != tbs:
                raise ValueError(f"ds_config: train_batch_size {tbs} != micro batch {self.micro_batch} x accumulation "
                                 f"steps {self.gas} x data-parallel size {dp_world}")
        self.stage = ps.pp_rank if ps else 0
        # pipeline bubble filling: split backward into dgrad and deferred wgrad (parallel/pipeline.py `schedule_zb`);
        # "pipeline": {"zero_bubble": false} restores plain 1F1B
        self.zero_bubble = pp_world > 1 and bool((config.get("pipeline") or {}).get("zero_bubble", True))
        if self.zero_bubble:
            from ..ops import gemm as _G

            _G.DEFER_CAPABLE = True

        fused_tp = None
        if tp_world > 1 and use_fused:
            from ..parallel.fused_tp import FusedTP

            fused_tp = FusedTP.create(ps.tp_group, ps.tp_rank, tp_world, self.device)
        self.tp = TPContext(ps.tp_group, tp_world, ps.tp_rank,
                            sequence_parallel=bool(config.get("sequence_parallel", True)),
                            fused=fused_tp) if tp_world > 1 else SINGLE

        if model_config.num_experts and pp_world > 1:
            # like the pinned Megatron-DeepSpeed: the load-balancing losses of earlier stages would be lost
            raise ValueError("MoE layers and pipeline parallelism cannot be combined (--num-experts needs "
                             "--pipeline-model-parallel-size 1)")
        # expert parallelism: EP groups are carved out of the DP dimension
        self.ep_world = ps.topo.ep if ps else 1
        self.ep_ctx = None
        if model_config.num_experts and self.ep_world > 1:
            from ..parallel.moe import EPContext

            ep_rank = ps.ep_ranks.index(ps.rank)
            comm = None
            if use_fused:
                from ..comm.arena import PeerComm

                comm = PeerComm(ps.ep_group, ep_rank, self.ep_world, self.device)
            self.ep_ctx = EPContext(ps.ep_group, self.ep_world, ep_rank, comm)
            if use_fused:       # K5: the quantised copies of the expert weights live until the next optimizer step
                from ..parallel.fused_moe import FusedMoE

                self.ep_ctx.fused_moe = FusedMoE(self.ep_ctx)
                self.ep_ctx.fused_moe.cache_weights = True
        # same init on every TP/DP replica of a stage, different across TP shards
        torch.manual_seed(seed + 1000 * self.stage + (ps.tp_rank if ps else 0))
        lo, hi = split_layers(model_config.layers, pp_world)[self.stage]
        self.model = GPTModel(model_config, self.tp, (lo, hi), pre=self.stage == 0,
                              post=self.stage == pp_world - 1, dtype=dtype, device=self.device,
                              ep=self.ep_ctx, fused_moe=use_fused)
        if model_config.num_experts and self.ep_world > 1:
            # every EP rank holds different experts: re-draw them with a rank-dependent seed
            g = torch.Generator(device=self.device).manual_seed(seed + 7919 * (self.ep_ctx.rank + 1))
            for n, p in self.model.named_parameters():
                if getattr(p, "is_expert", False) and p.dim() > 2:
                    p.data.normal_(0, model_config.init_std, generator=g)
        if tp_world > 1:
            # tensors every TP rank holds whole (embedding tables, norms, row-parallel biases) must start equal:
            # the per-rank seed above only suits the sharded weights
            for n, p in self.model.named_parameters():
                if n.startswith("embed.") or getattr(p, "tp_replicated", False):
                    dist.broadcast(p.data, src=ps.tp_ranks[0], group=ps.tp_group)
        # DeepSpeed-Ulysses (`--ds-sequence-parallel-size`): ranks of a group share each micro-batch and split its
        # sequence; outside attention everything is local, for gradients they are ordinary data-parallel ranks
        self.sp_world = ps.topo.sp if ps else 1
        if self.sp_world > 1:
            if tp_world > 1 or pp_world > 1 or model_config.num_experts:
                raise ValueError("sequence_parallel_size > 1 (Ulysses) is built for TP = PP = 1 dense models")
            for layer in self.model.layers:
                layer.ulysses_group = ps.sp_group
            self.model.seq_mult = self.sp_world
        self.model.recompute = bool(config.get("activation_checkpointing", {}).get("enabled", False))
        self.dp_comm = None
        allocator = None
        if dp_world > 1 and use_fused:
            from ..comm.arena import PeerComm

            self.dp_comm = PeerComm(ps.dp_group, ps.dp_rank, dp_world, self.device)
            allocator = self.dp_comm.allocator()
        is_exp = lambda n, p: getattr(p, "is_expert", False)          # noqa: E731
        self.flat = FlatParams(self.model, pad_to_multiple=max(1, dp_world) * 128, allocator=allocator,
                               param_filter=lambda n, p: not is_exp(n, p))
        has_experts = any(is_exp(n, p) for n, p in self.model.named_parameters())
        self.flat_expert = None
        if has_experts:
            edp = dp_world // self.ep_world
            self.flat_expert = FlatParams(self.model, pad_to_multiple=max(1, edp) * 128, param_filter=is_exp)
        if dp_world > 1:  # replicas start identical
            dist.broadcast(self.flat.params, src=ps.dp_ranks[0], group=ps.dp_group)

        op = config.get("optimizer", {}).get("params", {})
        betas = op.get("betas", (0.9, 0.95))
        acfg = AdamConfig(lr=float(op.get("lr", 1.5e-4)), beta1=float(betas[0]), beta2=float(betas[1]),
                          eps=float(op.get("eps", 1e-8)), weight_decay=float(op.get("weight_decay", 0.01)),
                          clip_grad=float(config.get("gradient_clipping", 1.0)))
        fused_dp = None
        if self.dp_comm is not None:
            from ..parallel.fused_dp import FusedZero1

            fused_dp = FusedZero1(self.dp_comm, self.flat.grads_handle, self.flat.params_handle)
        norm_groups = [ps.tp_group, ps.pp_group] if ps else []
        self.optimizer = Zero1Optimizer(self.flat, acfg, ps.dp_group if ps else None, dp_world,
                                        ps.dp_rank if ps else 0, norm_groups, fused=fused_dp)
        self.expert_optimizer = None
        if self.flat_expert is not None:
            edp = dp_world // self.ep_world
            eg = ps.expert_dp_group if (ps and self.ep_world > 1) else (ps.dp_group if ps else None)
            er = (ps.dp_rank // self.ep_world) if ps else 0
            extra = [ps.ep_group] if (ps and self.ep_world > 1) else []
            self.expert_optimizer = Zero1Optimizer(self.flat_expert, acfg, eg if edp > 1 else None, edp,
                                                   er if edp > 1 else 0, norm_groups + extra)
        self._sync_tied_weights()
        if self.pp_world > 1 and self.cfg.tie_embeddings and self.stage == pp_world - 1:
            o = self.optimizer      # master copy of the re-tied head rows
            o.master.copy_(self.flat.params[o.lo:o.hi].float())
        self.sched = config.get("scheduler", {}).get("params", None)
        from ..parallel.lr_schedules import make_schedule

        self._lr_fn = make_schedule(config.get("scheduler"), acfg.lr)     # Megatron's block or a DeepSpeed type
        # "fp16": {"enabled": true} keeps DeepSpeed's loss-scale state machine (parallel/loss_scaler.py)
        self.loss_scaler = make_loss_scaler(config.get("fp16"))
        self.skipped_steps = 0
        self._overflow = torch.zeros(1, dtype=torch.int32, device=self.device)
        # DeepSpeed's `wall_clock_breakdown` / `flops_profiler`: off unless the config asks
        self.timers = None
        if config.get("wall_clock_breakdown"):
            from ..obs.tracking import Timers

            self.timers = Timers(use_cuda=self.device.type == "cuda")
        self.comms_logger = None
        cl = config.get("comms_logger") or {}
        if cl.get("enabled"):       # DeepSpeed's comms_logger: library collectives and fused comm kernels per step
            from ..obs.comms import CommsLogger

            self.comms_logger = CommsLogger(sync=bool(cl.get("prof_all", False)), verbose=bool(cl.get("verbose", False))).enable()
        self.steps_per_print = int(config.get("steps_per_print", 10))
        fp = config.get("flops_profiler") or {}
        self._flops_profile_step = int(fp.get("profile_step", 1)) if fp.get("enabled") else -1
        self.global_steps = 0
        self._micro_steps, self._stepped_at = 0, -1
        self._drop_calls, self._seed = 0, int(seed)
        self.p2p = None
        self.last_loss = None

    # ------------------------------------------------------------- batch API
    def _p2p(self, seq: int):
        if self.p2p is None and self.pp_world > 1:
            tokens = seq * self.micro_batch
            if self.tp_world > 1 and self.tp.sequence_parallel:
                tokens //= self.tp_world
            ps = self.ps
            self.p2p = PP.P2P(ps.prev_stage_rank() if not ps.is_first_stage else None,
                              ps.next_stage_rank() if not ps.is_last_stage else None,
                              (tokens, self.cfg.hidden), self.dtype, self.device)
        return self.p2p

    def _set_drop_seed(self):
        """One dropout seed per micro-batch: equal on the TP / PP ranks of a replica, different between replicas;
        a recomputed forward (activation checkpointing) reads the same value again."""
        if (self.cfg.dropout > 0 or self.cfg.attention_dropout > 0) and self.model.training:
            self._drop_calls += 1
            self.model.drop_seed = (self._seed * 1000003 + self._drop_calls * 4099
                                    + (self.ps.dp_rank if self.ps else 0) * 15485863) & 0x3FFFFFFF

    def _sp_slice(self, tokens, labels):
        """This rank's slice of the sequence under Ulysses sequence parallelism (identity otherwise)."""
        if self.sp_world == 1:
            return tokens, labels
        assert tokens.shape[1] % self.sp_world == 0, "sequence length must divide by the sequence-parallel size"
        sl = tokens.shape[1] // self.sp_world
        lo = self.ps.sp_rank * sl
        self.model.embed.pos_offset = lo
        return tokens[:, lo:lo + sl].contiguous(), labels[:, lo:lo + sl].contiguous()

    def _sp_mean(self, loss):
        if self.sp_world > 1:  # mean over the whole sequence = mean of the slices' means
            loss = loss.clone()
            dist.all_reduce(loss, group=self.ps.sp_group)
            loss /= self.sp_world
        return loss

    def _forward_micro(self, batch, act):
        tokens, labels = self._sp_slice(*batch)
        b, s = tokens.shape
        last = self.stage == self.pp_world - 1
        self._set_drop_seed()
        if self.stage == 0:
            out = self.model(tokens=tokens, labels=labels if self.pp_world == 1 else None)
        else:
            out = self.model(hidden=act, labels=labels if last else None, seq=s, batch=b)
        if last and self.loss_scaler is not None and torch.is_grad_enabled():
            out = out * self.loss_scaler.scale     # backward sees the scaled loss
        return out

    def train_batch(self, data_iter: Iterator) -> torch.Tensor:
        """One optimizer step over `gradient_accumulation_steps` micro-batches.

        `data_iter` yields (tokens[b, s], labels[b, s]) device tensors; every pipeline stage
        of a DP replica must see the same micro-batches (the caller's loader guarantees it)."""
        self.optimizer.zero_grad()
        if self.expert_optimizer is not None:
            self.expert_optimizer.zero_grad()
        batches = [next(data_iter) for _ in range(self.gas)]
        tm = self.timers
        if tm is not None or self._flops_profile_step == self.global_steps + 1:
            import time

            self._sync()
            t_step = time.perf_counter()
        if self.pp_world == 1:
            losses = []
            for mb in batches:
                tm and tm.start("fwd")
                loss = self._forward_micro(mb, None)
                tm and (tm.stop("fwd"), tm.start("bwd"))
                loss.backward()
                tm and tm.stop("bwd")
                losses.append(loss.detach())
        else:
            seq = batches[0][0].shape[1]
            tm and tm.start("pipeline_fwd_bwd")
            losses = PP.run_1f1b(self.stage, self.pp_world, self.gas,
                                 lambda i, act: self._forward_micro(batches[i], act), self._p2p(seq),
                                 zero_bubble=self.zero_bubble)
            tm and tm.stop("pipeline_fwd_bwd")
        tm and tm.start("step")
        scale = self._optimizer_step()
        tm and tm.stop("step")
        if tm is not None and self.global_steps % self.steps_per_print == 0:
            ms = tm.summary(reset=True)                     # max over ranks
            if not dist.is_initialized() or dist.get_rank() == 0:
                print("time (ms) | " + " | ".join(f"{k}: {v / self.steps_per_print:.2f}" for k, v in sorted(ms.items())),
                      flush=True)
        if self.comms_logger is not None and self.global_steps % self.steps_per_print == 0:
            if not dist.is_initialized() or dist.get_rank() == 0:
                self.comms_logger.log_summary()
            self.comms_logger.reset()
        if self._flops_profile_step == self.global_steps:
            self._sync()
            self._print_flops_profile(batches, time.perf_counter() - t_step)
        loss = torch.stack(losses).mean() if losses else torch.zeros((), device=self.device)
        if scale != 1.0:
            loss = loss / scale
        if self.pp_world > 1:  # make the loss visible on every stage (for logging)
            dist.broadcast(loss, src=self.ps.pp_ranks[-1], group=self.ps.pp_group)
        loss = self._sp_mean(loss)
        self.last_loss = loss
        return loss

    def _sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize()

    def _print_flops_profile(self, batches, seconds: float) -> None:
        """DeepSpeed flops-profiler summary of one step: parameters, model FLOPs, achieved rate."""
        b, s = batches[0][0].shape
        tokens = b * s * self.gas * self.dp_world
        flops = self.cfg.flops_per_token(s) * tokens
        world = dist.get_world_size() if dist.is_initialized() else 1
        self.flops_profile = {"step": self.global_steps, "params": self.cfg.param_count(), "tokens_per_step": tokens,
                              "flops_per_step": flops, "seconds": seconds,
                              "tflops_per_gpu": flops / seconds / world / 1e12}
        if not dist.is_initialized() or dist.get_rank() == 0:
            p = self.flops_profile
            print(f"flops profiler | step {p['step']} | params {p['params'] / 1e6:.1f} M | fwd+bwd flops per step "
                  f"{p['flops_per_step'] / 1e12:.3f} T | step latency {seconds * 1e3:.1f} ms | "
                  f"{p['tflops_per_gpu']:.2f} TFLOPS per GPU", flush=True)

    def _optimizer_step(self) -> float:
        """Everything after the last backward of a step: gradient collection, TP / tied-weight reductions,
        fp16 overflow check, ZeRO-1 update.  Returns the loss scale the step ran with."""
        self.flat.collect_autograd_grads()
        if self.flat_expert is not None:
            self.flat_expert.collect_autograd_grads()
        allreduce_replicated_grads(self.flat, self.tp)
        if self.tp_world > 1:
            for _, p in self.flat.named_parameters():
                if getattr(p, "tp_sum_grad", False):
                    dist.all_reduce(p.main_grad, group=self.ps.tp_group)
        self._reduce_tied_grads()
        div, scale = float(self.gas * self.dp_world), 1.0
        overflow = False
        if self.loss_scaler is not None:
            scale = self.loss_scaler.scale
            div *= scale                            # the unscale rides on the gradient divisor
            overflow = self._grads_overflowed()
            self.loss_scaler.update(overflow)
        if overflow:                                # DeepSpeed: skip the update and the lr step
            self.skipped_steps += 1
        else:
            self._apply_lr()
            self.optimizer.step(grad_divisor=div)
            if self.expert_optimizer is not None:
                self.expert_optimizer.lr = self.optimizer.lr
                self.expert_optimizer.step(grad_divisor=div)
        self.global_steps += 1
        fm = getattr(self.ep_ctx, "fused_moe", None) if self.ep_ctx is not None else None
        if fm is not None:
            fm.invalidate()                          # the expert weights just changed
        return scale

    # ------------------------------------------- DeepSpeed's forward / backward / step triple
    def __call__(self, tokens: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return self.forward(tokens, labels)

    def forward(self, tokens: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """`loss = engine(batch)` of a DeepSpeed training loop without pipeline parallelism (Megatron-DeepSpeed
        takes this route when `--pipeline-model-parallel-size 1`: `model[0](...)`, `model[0].backward(loss)`,
        `model[0].step()`).  Returns the unscaled loss of one micro-batch."""
        assert self.pp_world == 1, "forward/backward/step is the non-pipeline API; use train_batch under PP"
        if self._micro_steps % self.gas == 0:
            self.optimizer.zero_grad()
            if self.expert_optimizer is not None:
                self.expert_optimizer.zero_grad()
        self._set_drop_seed()
        tokens, labels = self._sp_slice(tokens, labels)
        return self.model(tokens=tokens, labels=labels)

    def backward(self, loss: torch.Tensor) -> torch.Tensor:
        (loss * self.loss_scaler.scale if self.loss_scaler is not None else loss).backward()
        self._micro_steps += 1
        return loss

    def is_gradient_accumulation_boundary(self) -> bool:
        return self._micro_steps % self.gas == 0 and self._micro_steps > 0

    def step(self) -> None:
        """Optimizer step at a gradient-accumulation boundary, a no-op otherwise (DeepSpeed semantics)."""
        if not self.is_gradient_accumulation_boundary() or self._stepped_at == self._micro_steps:
            return
        self._stepped_at = self._micro_steps
        self._optimizer_step()

    def eval_batch(self, data_iter: Iterator, num_micro_batches: int = 1) -> torch.Tensor:
        """Forward-only loss over `num_micro_batches` micro-batches (Megatron's `--eval-iters` loop calls
        this; examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml:57).  Under pipeline
        parallelism the micro-batches stream through the stages and every stage returns the loss."""
        batches = [next(data_iter) for _ in range(num_micro_batches)]
        was_training = self.model.training
        self.model.eval()                  # no dropout while evaluating
        try:
            return self._eval_micro_batches(batches, num_micro_batches)
        finally:
            self.model.train(was_training)

    def _eval_micro_batches(self, batches, num_micro_batches):
        with torch.no_grad():
            if self.pp_world == 1:
                losses = [self.model(tokens=t, labels=l) for t, l in (self._sp_slice(*b) for b in batches)]
            else:
                seq = batches[0][0].shape[1]
                losses = PP.run_forward_only(self.stage, self.pp_world, num_micro_batches,
                                             lambda i, act: self._forward_micro(batches[i], act), self._p2p(seq))
            loss = torch.stack(losses).mean() if losses else torch.zeros((), device=self.device)
            if self.pp_world > 1:
                dist.broadcast(loss, src=self.ps.pp_ranks[-1], group=self.ps.pp_group)
        return self._sp_mean(loss)

    def _tied_slice(self, t):
        """Stage 0 keeps the whole (replicated) embedding table, the last stage a vocab-parallel
        head: this TP rank's rows of the table are the ones tied to its head shard."""
        if self.stage != 0 or self.tp_world == 1:
            return t
        n = t.shape[0] // self.tp_world
        return t[self.ps.tp_rank * n:(self.ps.tp_rank + 1) * n]

    def _sync_tied_weights(self):
        """Make the last stage's head start as a copy of the first stage's embedding rows."""
        ps = self.ps
        if ps is None or self.pp_world == 1 or not self.cfg.tie_embeddings or ps.embedding_group is None:
            return
        w = self.model.embed.wte if self.stage == 0 else self.model.head.weight
        buf = self._tied_slice(w.data).contiguous()
        dist.broadcast(buf, src=ps.pp_ranks[0], group=ps.embedding_group)
        if self.stage != 0:
            w.data.copy_(buf)

    def _reduce_tied_grads(self):
        ps = self.ps
        if ps is None or self.pp_world == 1 or not self.cfg.tie_embeddings:
            return
        if ps.embedding_group is not None:
            w = self.model.embed.wte if self.stage == 0 else self.model.head.weight
            g = self._tied_slice(w.main_grad)
            dist.all_reduce(g, group=ps.embedding_group)
            if self.stage == 0 and self.tp_world > 1:
                # every TP replica of the table needs every slice's tied gradient
                dist.all_gather_into_tensor(w.main_grad, g.clone(), group=ps.tp_group) if g.is_cuda else \
                    self._cpu_gather_rows(w.main_grad, g)

    def _cpu_gather_rows(self, full, mine):
        parts = [torch.empty_like(mine) for _ in range(self.tp_world)]
        dist.all_gather(parts, mine.contiguous(), group=self.ps.tp_group)
        full.copy_(torch.cat(parts, 0))

    def _grads_overflowed(self) -> bool:
        """inf/nan anywhere in this rank's gradients, agreed on by every rank (MAX all-reduce:
        SURVEY.md §2H C6).  Costs one pass over the gradients and a host sync, fp16 mode only."""
        from ..ops import optim as O

        self._overflow.zero_()
        for flat in (self.flat, self.flat_expert):
            if flat is None:
                continue
            if flat.grads.is_cuda:
                O.sumsq(flat.grads, overflow=self._overflow)
            elif not bool(torch.isfinite(flat.grads).all()):
                self._overflow.fill_(1)
        if self.ps is not None and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self._overflow, op=dist.ReduceOp.MAX)
        return bool(self._overflow.item())

    @property
    def loss_scale(self) -> float:
        return self.loss_scaler.scale if self.loss_scaler is not None else 1.0

    def _apply_lr(self):
        if self._lr_fn is not None:
            self.optimizer.lr = self._lr_fn(self.global_steps - self.skipped_steps + 1)

    # ----------------------------------------------------------- checkpoints
    def save_checkpoint(self, save_dir: str, tag: str | None = None, client_state: dict | None = None,
                        save_optimizer: bool = True, save_rng: bool = True, async_save: bool = False):
        """`async_save=True` (or `"checkpoint": {"async_save": true}` in the config) returns once tensors are staged on the
        host; files and `latest` follow on background threads (`ckpt.megatron_ds_layout.wait_for_pending_saves`)."""
        from ..ckpt.megatron_ds_layout import save_checkpoint

        async_save = async_save or bool((self.ds.get("checkpoint") or {}).get("async_save", False))
        return save_checkpoint(self, save_dir, tag, client_state or {}, async_save=async_save,
                               save_optimizer=save_optimizer, save_rng=save_rng)

    def load_checkpoint(self, load_dir: str, tag: str | None = None, load_optimizer_states: bool = True,
                        load_rng: bool = True, finetune: bool = False):
        from ..ckpt.megatron_ds_layout import load_checkpoint

        return load_checkpoint(self, load_dir, tag, load_optimizer_states, load_rng, finetune)

    # ----------------------------------------------------- DeepSpeed accessors
    @property
    def module(self):
        return self.model

    def get_lr(self) -> list[float]:
        return [self.optimizer.lr]

    def get_global_grad_norm(self) -> float:
        return self.grad_norm()

    def train_micro_batch_size_per_gpu(self) -> int:
        return self.micro_batch

    def gradient_accumulation_steps(self) -> int:
        return self.gas

    def train_batch_size(self) -> int:
        return self.micro_batch * self.gas * (self.dp_world // self.sp_world)

    def train(self, mode: bool = True):
        self.model.train(mode)
        return self

    def eval(self):
        return self.train(False)

    def grad_norm(self) -> float:
        g = self.optimizer.last_grad_norm
        return float(g.item()) if g is not None else float("nan")


def initialize(model_config: GPTConfig | None = None, config: dict | None = None,
               parallel: ParallelState | None = None, **kw) -> TrainEngine:
    """DeepSpeed-style entry point (`deepspeed.initialize`)."""
    return TrainEngine(model_config, config or {}, parallel, **kw)


def initialize_tuple(model_config: GPTConfig | None = None, config: dict | None = None,
                     parallel: ParallelState | None = None, **kw):
    """`engine, optimizer, dataloader, lr_scheduler = deepspeed.initialize(...)`: the four-tuple form scripts
    unpack (the data loader is the caller's; the schedule lives in the engine)."""
    eng = initialize(model_config, config, parallel, **kw)
    return eng, eng.optimizer, None, eng
