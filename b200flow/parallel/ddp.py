"""Data-parallel wrapper with a bucketed gradient reducer (K1).

What torch's C++ `Reducer` (25 MB buckets in reverse registration order, one all-reduce per
bucket as soon as its gradients exist, overlapped with the rest of backward) and Horovod's
fusion buffer do for the reference's DDP / Horovod workloads (SURVEY.md §2F "DP", §2H C1/C2),
re-designed around flat storage:

* gradients live in ONE flat fp32 buffer (symmetric peer memory on GPUs), so a bucket is a
  slice and needs no copy-in / copy-out;
* buckets are launched strictly in index order (a complete bucket waits behind an incomplete earlier one), so
  every rank issues the same sequence of reductions whatever order its autograd hooks fire in;
* the bucket plan comes from the host C++ planner (`csrc/host/bucket_plan.cpp`) and the per-event
  readiness bookkeeping runs in the host C++ reducer (`csrc/host/reducer.cpp`): a gradient-ready event is
  one call that answers "which bucket, if any, is now complete";
* a bucket is reduced by the K1 kernel (`csrc/comm/collectives.cu`): in-place NVLS /
  peer-memory all-reduce with the 1/N averaging fused, launched on a side stream using a
  few SMs while backward keeps the rest;
* readiness is tracked per parameter: autograd post-accumulate hooks for gradients autograd
  produces, and the `grad_ready` callback that the wgrad GEMM / norm kernels fire after
  accumulating straight into `main_grad`.

On CPU (gloo) the same bookkeeping drives `dist.all_reduce` on the slice.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch
import torch.distributed as dist
from torch import nn

from ..utils.flat import FlatParams


def plan_buckets(flat: FlatParams, cap_mb: float = 25.0, first_cap_mb: float = 1.0) -> list[tuple]:
    """[(first_slot, last_slot_excl, elem_begin, elem_end)] in backward-fill order."""
    from ..ops._lib import host_lib

    lib = host_lib()
    n = len(flat.slots)
    offs = np.array([s.offset for s in flat.slots], dtype=np.int64)
    nums = np.array([s.numel for s in flat.slots], dtype=np.int64)
    out = np.zeros((n + 1) * 4, dtype=np.int64)
    lib.b2f_plan_buckets.restype = ctypes.c_int
    nb = lib.b2f_plan_buckets(offs.ctypes.data_as(ctypes.c_void_p), nums.ctypes.data_as(ctypes.c_void_p),
                              ctypes.c_int(n), ctypes.c_longlong(flat.numel),
                              ctypes.c_longlong(flat.grads.element_size()),
                              ctypes.c_longlong(int(cap_mb * (1 << 20))),
                              ctypes.c_longlong(int(first_cap_mb * (1 << 20))),
                              out.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n + 1))
    assert nb > 0, "bucket planner failed"
    return [tuple(int(v) for v in out[4 * i:4 * i + 4]) for i in range(nb)]


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, process_group=None, bucket_cap_mb: float = 25.0,
                 fused: bool = True, comm_blocks: int = 32):
        super().__init__()
        self.module = module
        self.group = process_group if process_group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        p0 = next(module.parameters())
        self.on_gpu = p0.is_cuda
        self.comm = None
        allocator = None
        if self.on_gpu and fused and self.world > 1:
            from ..comm.arena import PeerComm

            self.comm = PeerComm(self.group, self.rank, self.world, p0.device)
            allocator = self.comm.allocator()
        self.flat = FlatParams(module, allocator=allocator, pad_to_multiple=128)
        if self.world > 1:
            dist.broadcast(self.flat.params, src=dist.get_global_rank(self.group, 0) if hasattr(
                dist, "get_global_rank") else 0, group=self.group)
        self.require_sync = True
        self.timeline = None          # optional obs.tracking.Timeline: one event per bucket reduction
        self.reducer = None
        self.side = None
        if self.comm is not None:
            from .fused_dp import FusedAllReduce

            self.reducer = FusedAllReduce(self.comm, self.flat.grads_handle, blocks=comm_blocks)
            self.side = torch.cuda.Stream(device=p0.device)
        self._works = []
        params = dict(self.flat.named_parameters())
        self._params_by_slot = [params[slot.name] for slot in self.flat.slots]
        for si, slot in enumerate(self.flat.slots):
            p = params[slot.name]
            p.grad_ready = self._make_ready(si)
            p.register_post_accumulate_grad_hook(self._make_hook(si))
        self.set_bucket_cap(bucket_cap_mb)

    # ------------------------------------------------------------- bookkeeping
    def set_bucket_cap(self, cap_mb: float) -> None:
        """(Re)plan the buckets.  Only between steps (after `finish_reduction`); every rank must pass the
        same value — the plan is a pure function of the parameter layout and the cap."""
        from ..ops._lib import host_lib

        self.bucket_cap_mb = float(cap_mb)
        self.buckets = plan_buckets(self.flat, cap_mb, first_cap_mb=min(1.0, cap_mb))
        lib = host_lib()
        lib.b2f_reducer_create.restype = ctypes.c_void_p
        lib.b2f_reducer_mark.restype = lib.b2f_reducer_next_unlaunched.restype = ctypes.c_int
        lib.b2f_reducer_next_ready.restype = ctypes.c_int
        lib.b2f_reducer_mark.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.b2f_reducer_next_unlaunched.argtypes = lib.b2f_reducer_reset.argtypes = [ctypes.c_void_p]
        lib.b2f_reducer_next_ready.argtypes = [ctypes.c_void_p]
        lib.b2f_reducer_destroy.argtypes = [ctypes.c_void_p]
        self._free_reducer()
        plan = np.array(self.buckets, dtype=np.int64).reshape(-1)
        uses = np.array([int(getattr(p, "ddp_uses", 1)) for p in self._params_by_slot], dtype=np.int32)
        self._lib = lib
        self._state = lib.b2f_reducer_create(plan.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(self.buckets)),
                                             ctypes.c_int(len(self.flat.slots)), uses.ctypes.data_as(ctypes.c_void_p))
        assert self._state, "bucket plan does not cover every parameter exactly once"
        self._uses_seen = uses.tolist()

    def _free_reducer(self):
        if getattr(self, "_state", None):
            self._lib.b2f_reducer_destroy(self._state)
            self._state = None

    def __del__(self):
        try:
            self._free_reducer()
        except Exception:       # noqa: BLE001 - interpreter shutdown
            pass

    def _reset(self):
        uses = [int(getattr(p, "ddp_uses", 1)) for p in self._params_by_slot]
        if uses != self._uses_seen:          # a model changed how often it uses a parameter: rebuild
            self.set_bucket_cap(self.bucket_cap_mb)
        else:
            self._lib.b2f_reducer_reset(self._state)

    def _make_hook(self, si):
        def hook(p):
            if p.grad is not None:
                p.main_grad.add_(p.grad.to(p.main_grad.dtype))
                p.grad = None
            self._mark(si)
        return hook

    def _make_ready(self, si):
        return lambda: self._mark(si)

    def _mark(self, si):
        if not self.require_sync:
            return
        bi = self._lib.b2f_reducer_mark(self._state, si)
        while bi >= 0:          # buckets leave in index order; one event can release several in a row
            self._launch(bi)
            bi = self._lib.b2f_reducer_next_ready(self._state)

    def _launch(self, bi):
        if self.world == 1:
            return
        _, _, e0, e1 = self.buckets[bi]
        if self.timeline is not None:
            self.timeline.instant("allreduce", bucket=bi, bytes=(e1 - e0) * self.flat.grads.element_size())
        if self.reducer is not None:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                self.reducer(e0, e1 - e0, 1.0 / self.world)
        else:
            g = self.flat.grads[e0:e1]
            if g.is_cuda:
                self._works.append((dist.all_reduce(g, group=self.group, async_op=True), g))
            else:
                dist.all_reduce(g, group=self.group)
                g.div_(self.world)

    # ----------------------------------------------------------------- public
    def forward(self, *a, **kw):
        return self.module(*a, **kw)

    def no_sync(self):
        ddp = self

        class _Ctx:
            def __enter__(self):
                ddp.require_sync = False

            def __exit__(self, *exc):
                ddp.require_sync = True
        return _Ctx()

    def finish_reduction(self):
        """Call after backward: reduce any bucket whose parameters received no gradient this
        step, wait for the side stream, re-arm the bookkeeping."""
        if self.require_sync:
            while (bi := self._lib.b2f_reducer_next_unlaunched(self._state)) >= 0:
                self._launch(bi)
            if self.side is not None:
                torch.cuda.current_stream().wait_stream(self.side)
            for w, g in self._works:
                w.wait()
                g.div_(self.world)
            self._works = []
        self._reset()

    def zero_grad(self, set_to_none: bool = True):
        self.flat.zero_grad()


class BucketAutotuner:
    """Picks the bucket (fusion-buffer) size by measurement, in the spirit of `HOROVOD_AUTOTUNE=1`
    — which the reference's MPI chart switches on by default
    (charts/machine-learning/training/mpijob-horovod-tensorflow-gpu/values.yaml:22-24,38-39).

    Horovod scores (fusion threshold, cycle time) pairs by bytes reduced per second; here the only
    knob is the bucket cap, and the score is the step time itself.  After `warmup` steps each
    candidate runs for `steps_per_sample` steps; the per-candidate time is the MAX over ranks (one
    tiny all-reduce per sample), so every rank takes the same decision; the fastest cap is then kept.
    Call `step_done()` once per training step, after `finish_reduction()` / the optimizer step."""

    def __init__(self, ddp: DistributedDataParallel, candidates_mb=(1.0, 4.0, 8.0, 16.0, 25.0, 64.0, 128.0),
                 warmup: int = 3, steps_per_sample: int = 10, log_path: str | None = None, clock=None):
        import time

        self.ddp, self.cands = ddp, [float(c) for c in candidates_mb]
        self.warmup, self.sps, self.log_path = warmup, max(1, steps_per_sample), log_path
        self.clock = clock or time.perf_counter
        self.results: list[tuple[float, float]] = []     # (cap_mb, seconds per step)
        self.done = ddp.world == 1 or not self.cands
        self.best_mb = ddp.bucket_cap_mb
        self._i, self._n, self._t0 = -1, 0, None          # candidate index (-1 = warm-up), steps seen, start

    def _sync_time(self) -> float:
        if self.ddp.on_gpu:
            torch.cuda.synchronize()
        return self.clock()

    def step_done(self) -> None:
        if self.done:
            return
        self._n += 1
        if self._i < 0:
            if self._n >= self.warmup:
                self._start(0)
            return
        if self._n < self.sps:
            return
        dt = (self._sync_time() - self._t0) / self.sps
        t = torch.tensor([dt], dtype=torch.float64, device=self.ddp.flat.grads.device if self.ddp.on_gpu else "cpu")
        if self.ddp.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.ddp.group)
        self.results.append((self.cands[self._i], float(t)))
        if self._i + 1 < len(self.cands):
            self._start(self._i + 1)
            return
        self.best_mb = min(self.results, key=lambda r: r[1])[0]
        self.ddp.set_bucket_cap(self.best_mb)
        self.done = True
        if self.log_path and self.ddp.rank == 0:
            with open(self.log_path, "w") as f:        # Horovod's HOROVOD_AUTOTUNE_LOG is a CSV too
                f.write("bucket_cap_mb,seconds_per_step,chosen\n")
                for cap, sec in self.results:
                    f.write(f"{cap},{sec:.6f},{int(cap == self.best_mb)}\n")

    def _start(self, i: int) -> None:
        self._i, self._n = i, 0
        self.ddp.set_bucket_cap(self.cands[i])
        self._t0 = self._sync_time()
