"""Horovod-shaped facade (`import b200flow.engines.horovod_api as hvd`).

The reference's Mask R-CNN workloads run under Horovod 0.28.1 with NCCL GPU operations, launched
by `mpirun` (charts/machine-learning/training/mpijob-horovod-tensorflow-gpu/values.yaml:18-60,
containers/aws-samples-maskrcnn/Dockerfile:57).  What a training script touches is small: `init`,
`rank / size / local_rank`, `allreduce`, `broadcast_parameters`, `DistributedOptimizer` — and
underneath, the tensor-fusion buffer that batches many small gradients into few collectives.

Here the ranks come from the launcher's mpirun mode (`OMPI_COMM_WORLD_*`, or torchrun's variables),
collectives are `torch.distributed` (NCCL on GPUs, gloo on CPU), and the fusion buffer is the flat
gradient storage + bucketed K1 reducer of `parallel/ddp.py` (gradients are *born* inside one
contiguous buffer, so "fusion" needs no copy-in / copy-out).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

Average, Sum = "average", "sum"


def init(backend: str | None = None) -> None:
    if dist.is_initialized():
        return
    env = os.environ
    rank = int(env.get("OMPI_COMM_WORLD_RANK", env.get("RANK", 0)))
    world = int(env.get("OMPI_COMM_WORLD_SIZE", env.get("WORLD_SIZE", 1)))
    local = int(env.get("OMPI_COMM_WORLD_LOCAL_RANK", env.get("LOCAL_RANK", 0)))
    env.setdefault("RANK", str(rank))
    env.setdefault("WORLD_SIZE", str(world))
    env.setdefault("LOCAL_RANK", str(local))
    env.setdefault("MASTER_ADDR", "127.0.0.1")
    env.setdefault("MASTER_PORT", "29500")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())
    if world > 1:
        dist.init_process_group(backend, rank=rank, world_size=world)


def shutdown() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()


def is_initialized() -> bool:
    return dist.is_initialized() or int(os.environ.get("WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_SIZE", 1))) == 1


def size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def local_rank() -> int:
    return int(os.environ.get("OMPI_COMM_WORLD_LOCAL_RANK", os.environ.get("LOCAL_RANK", 0)))


def local_size() -> int:
    return int(os.environ.get("OMPI_COMM_WORLD_LOCAL_SIZE", os.environ.get("LOCAL_WORLD_SIZE", size())))


def allreduce(tensor: torch.Tensor, op: str = Average, name: str | None = None) -> torch.Tensor:
    out = tensor.clone()
    if size() > 1:
        dist.all_reduce(out)
        if op == Average:
            out = out / size() if out.is_floating_point() else out // size()
    return out


def allgather(tensor: torch.Tensor, name: str | None = None) -> torch.Tensor:
    if size() == 1:
        return tensor.clone()
    outs = [torch.empty_like(tensor) for _ in range(size())]
    dist.all_gather(outs, tensor.contiguous())
    return torch.cat(outs, 0)


def broadcast(tensor: torch.Tensor, root_rank: int = 0, name: str | None = None) -> torch.Tensor:
    out = tensor.clone()
    if size() > 1:
        dist.broadcast(out, src=root_rank)
    return out


def allreduce_(tensor: torch.Tensor, op: str = Average, name: str | None = None) -> torch.Tensor:
    """In-place form of `allreduce`."""
    tensor.copy_(allreduce(tensor, op))
    return tensor


def alltoall(tensor: torch.Tensor, splits=None, name: str | None = None) -> torch.Tensor:
    """Equal splits along dim 0: chunk r goes to rank r (gloo: through an all-gather)."""
    n = size()
    if n == 1:
        return tensor.clone()
    assert splits is None and tensor.shape[0] % n == 0, "alltoall: equal splits along dim 0"
    if tensor.is_cuda:
        out = torch.empty_like(tensor)
        dist.all_to_all_single(out, tensor.contiguous())
        return out
    everyone = [torch.empty_like(tensor) for _ in range(n)]
    dist.all_gather(everyone, tensor.contiguous())
    k = tensor.shape[0] // n
    return torch.cat([everyone[r][rank() * k:(rank() + 1) * k] for r in range(n)], 0)


def reducescatter(tensor: torch.Tensor, op: str = Average, name: str | None = None) -> torch.Tensor:
    """Rank r keeps rows r of the reduced tensor (dim 0 split evenly)."""
    n = size()
    if n == 1:
        return tensor.clone()
    full = allreduce(tensor, op)
    k = tensor.shape[0] // n
    return full[rank() * k:(rank() + 1) * k].clone()


def barrier() -> None:
    if size() > 1:
        dist.barrier()


def join() -> int:
    """Ranks that ran out of data wait here for the others; returns the last rank to join (here: the last rank)."""
    barrier()
    return size() - 1


def broadcast_object(obj, root_rank: int = 0, name: str | None = None):
    if size() == 1:
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=root_rank)
    return box[0]


def allgather_object(obj, name: str | None = None) -> list:
    if size() == 1:
        return [obj]
    out = [None] * size()
    dist.all_gather_object(out, obj)
    return out


def cross_rank() -> int:
    """Node index (one box: always 0) and number of nodes, as Horovod reports them."""
    return 0


def cross_size() -> int:
    return 1


def nccl_built() -> bool:
    return torch.cuda.is_available()


def mpi_built() -> bool:
    return False            # ranks come from the launcher's mpirun mode, there is no MPI control plane


def gloo_built() -> bool:
    return True


def is_homogeneous() -> bool:
    return True


class Compression:
    """`hvd.Compression.none / fp16`: gradient compression for the all-reduce.  The flat fp32 gradient buffer is
    reduced in place by K1, so `fp16` is accepted and means the same as `none`."""
    none = "none"
    fp16 = "fp16"


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """`hvd.broadcast_parameters(model.state_dict(), root_rank=0)` — start every rank from rank 0's weights."""
    if size() == 1:
        return
    items = params.items() if hasattr(params, "items") else params
    for _, p in items:
        if isinstance(p, torch.Tensor):
            dist.broadcast(p.data if isinstance(p, torch.nn.Parameter) else p, src=root_rank)


def broadcast_optimizer_state(optimizer, root_rank: int = 0) -> None:
    if size() == 1:
        return
    for st in optimizer.state.values():
        for v in st.values():
            if isinstance(v, torch.Tensor):
                dist.broadcast(v, src=root_rank)


class _DistributedOptimizer:
    """Wraps an optimizer: gradients are averaged across ranks before `step()`.  With a model wrapped by
    `b200flow.parallel.ddp.DistributedDataParallel` the bucketed reducer has already done it during
    backward (pass `ddp=`); otherwise every gradient is all-reduced here (one flat collective per dtype:
    the fusion buffer)."""

    def __init__(self, optimizer, named_parameters=None, ddp=None, op: str = Average,
                 backward_passes_per_step: int = 1):
        self.optimizer, self.ddp, self.op = optimizer, ddp, op
        self.passes, self._seen = max(1, int(backward_passes_per_step)), 0
        self.params = [p for _, p in named_parameters] if named_parameters is not None else [
            p for g in optimizer.param_groups for p in g["params"]]

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    def synchronize(self) -> None:
        if size() == 1:
            return
        if self.ddp is not None:
            self.ddp.finish_reduction()
            return
        by_dtype: dict = {}
        for p in self.params:
            if p.grad is not None:
                by_dtype.setdefault(p.grad.dtype, []).append(p.grad)
        for grads in by_dtype.values():
            flat = torch.cat([g.reshape(-1) for g in grads])
            dist.all_reduce(flat)
            if self.op == Average:
                flat /= size()
            off = 0
            for g in grads:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()

    def step(self, closure=None):
        """With `backward_passes_per_step` = k the gradients of k backward passes are accumulated locally and
        reduced once: only every k-th call reduces and updates."""
        self._seen += 1
        if self._seen % self.passes:
            return None
        self.synchronize()
        return self.optimizer.step(closure) if closure is not None else self.optimizer.step()

    def zero_grad(self, *a, **kw):
        if self._seen % self.passes:
            return None            # inside an accumulation window the gradients are kept
        return self.optimizer.zero_grad(*a, **kw)

    def state_dict(self):
        return self.optimizer.state_dict()

    def load_state_dict(self, sd):
        return self.optimizer.load_state_dict(sd)


def ddp_options_from_env(env=None) -> dict:
    """Horovod's tuning knobs as the chart exports them (`mpirun -x HOROVOD_AUTOTUNE ...`,
    charts/machine-learning/training/mpijob-horovod-tensorflow-gpu/values.yaml:22-24,38-41,79-82):

    HOROVOD_FUSION_THRESHOLD  bytes of the fusion buffer   -> bucket cap of the reducer (default 64 MiB)
    HOROVOD_AUTOTUNE          "1": pick the cap by measurement (parallel/ddp.py BucketAutotuner)
    HOROVOD_AUTOTUNE_LOG      CSV of the candidates tried
    HOROVOD_AUTOTUNE_WARMUP_SAMPLES / HOROVOD_AUTOTUNE_STEPS_PER_SAMPLE
    HOROVOD_TIMELINE          chrome-trace file of the bucket reductions (rank 0)
    HOROVOD_CYCLE_TIME, HOROVOD_HIERARCHICAL_*, HOROVOD_TORUS_ALLREDUCE have no meaning without a
    background negotiation thread / multi-node topology and are accepted silently."""
    env = os.environ if env is None else env
    return {"bucket_cap_mb": int(env.get("HOROVOD_FUSION_THRESHOLD", 64 << 20)) / float(1 << 20),
            "autotune": env.get("HOROVOD_AUTOTUNE", "0") not in ("0", "", "false", "False"),
            "autotune_log": env.get("HOROVOD_AUTOTUNE_LOG") or None,
            "autotune_warmup": int(env.get("HOROVOD_AUTOTUNE_WARMUP_SAMPLES", 3)),
            "autotune_steps": int(env.get("HOROVOD_AUTOTUNE_STEPS_PER_SAMPLE", 10)),
            "timeline": env.get("HOROVOD_TIMELINE") or None}


def make_ddp(module, env=None, **kw):
    """(ddp, autotuner | None): the bucketed reducer configured from the Horovod environment."""
    from ..obs.tracking import Timeline
    from ..parallel.ddp import BucketAutotuner, DistributedDataParallel

    o = ddp_options_from_env(env)
    ddp = DistributedDataParallel(module, bucket_cap_mb=max(o["bucket_cap_mb"], 1.0 / 1024), **kw)
    if o["timeline"] and ddp.rank == 0:
        ddp.timeline = Timeline(o["timeline"], ddp.rank)
    tuner = None
    if o["autotune"] and ddp.world > 1:
        tuner = BucketAutotuner(ddp, warmup=o["autotune_warmup"], steps_per_sample=o["autotune_steps"],
                                log_path=o["autotune_log"])
        if ddp.rank == 0 and os.environ.get("HOROVOD_LOG_LEVEL", "").upper() in ("INFO", "DEBUG", "TRACE"):
            print(f"[b200flow] HOROVOD_AUTOTUNE: trying bucket caps {tuner.cands} MB", flush=True)
    return ddp, tuner


def DistributedOptimizer(optimizer, named_parameters=None, ddp=None, op: str = Average,
                         backward_passes_per_step: int = 1, compression=Compression.none, **_ignored):
    return _DistributedOptimizer(optimizer, named_parameters, ddp, op, backward_passes_per_step)
