"""Communication log: which collectives ran, how often, how many bytes, how long.

DeepSpeed's `comms_logger` (`"comms_logger": {"enabled": true, "verbose": ..., "prof_all": ...}` and
`deepspeed.comm.log_summary()`; SURVEY.md §5.1 lists it among the facilities of the pinned stack) records
every `torch.distributed` call it issues.  Here there are two kinds of communication to account for:

* library collectives (`torch.distributed`: NCCL on GPUs, gloo on CPU) — recorded by wrapping the
  functions of the module while the logger is enabled;
* the fused kernels, which move data themselves over NVLink peer memory (K1 all-reduce, K2 ZeRO-1 pair,
  K3 / K4 / K4' tensor-parallel GEMMs, K5 MoE dispatch / combine, K9 statistics exchange) — recorded by name
  through the native binding's call hook, since no library call exists to intercept.

Latency is wall-clock around the call; with `sync=True` the device is synchronised on both sides so the
number is the collective's own duration (DeepSpeed does the same when profiling), at the cost of
serialising the step.  Bus bandwidth uses the usual factors (all-reduce 2(n-1)/n, all-gather and
reduce-scatter (n-1)/n).
"""
from __future__ import annotations

import time
from collections import defaultdict

import torch
import torch.distributed as dist

_FUSED = {"b2f_allreduce_f32": "fused all-reduce (K1)", "b2f_zero1_reduce": "fused ZeRO-1 reduce-scatter (K2)",
          "b2f_zero1_adam_gather": "fused Adam + all-gather (K2)", "b2f_ag_gemm_bf16": "fused all-gather -> GEMM (K3)",
          "b2f_gemm_rs_bf16": "fused GEMM -> reduce-scatter (K4)", "b2f_gemm_ar_bf16": "fused GEMM -> all-reduce (K4')",
          "b2f_moe_dispatch": "MoE dispatch (K5)", "b2f_moe_combine": "MoE combine (K5)",
          "b2f_vocab_ce_stats": "vocab-parallel CE exchange (K9)", "b2f_peer_barrier": "peer flag barrier"}

_PATCHED = ("all_reduce", "all_gather", "all_gather_into_tensor", "reduce_scatter_tensor", "broadcast",
            "all_to_all_single", "reduce", "barrier", "send", "recv", "batch_isend_irecv")


def _nbytes(args, kwargs) -> int:
    total = 0
    for a in list(args) + list(kwargs.values()):
        if isinstance(a, torch.Tensor):
            total = max(total, a.numel() * a.element_size())
        elif isinstance(a, (list, tuple)):
            sub = 0
            for x in a:
                if isinstance(x, torch.Tensor):
                    sub += x.numel() * x.element_size()
                elif hasattr(x, "tensor") and isinstance(x.tensor, torch.Tensor):     # dist.P2POp
                    sub += x.tensor.numel() * x.tensor.element_size()
            total = max(total, sub)
    return total


def _bus_factor(op: str, n: int) -> float:
    if n <= 1:
        return 1.0
    if op == "all_reduce":
        return 2.0 * (n - 1) / n
    if op in ("all_gather", "all_gather_into_tensor", "reduce_scatter_tensor", "all_to_all_single"):
        return (n - 1) / n
    return 1.0


class CommsLogger:
    def __init__(self, sync: bool = False, verbose: bool = False):
        self.sync, self.verbose = sync, verbose
        self.records: dict = defaultdict(lambda: {"count": 0, "bytes": 0, "seconds": 0.0, "bus_bytes": 0.0})
        self.fused: dict = defaultdict(int)
        self._orig: dict = {}
        self.enabled = False

    # ------------------------------------------------------------------ switch
    def enable(self) -> "CommsLogger":
        if self.enabled:
            return self
        for name in _PATCHED:
            fn = getattr(dist, name, None)
            if fn is not None:
                self._orig[name] = fn
                setattr(dist, name, self._wrap(name, fn))
        from ..ops import _lib

        self._prev_hook = getattr(_lib, "call_hook", None)
        _lib.call_hook = self._on_native
        self.enabled = True
        return self

    def disable(self) -> None:
        if not self.enabled:
            return
        for name, fn in self._orig.items():
            setattr(dist, name, fn)
        self._orig = {}
        from ..ops import _lib

        _lib.call_hook = self._prev_hook
        self.enabled = False

    def __enter__(self):
        return self.enable()

    def __exit__(self, *exc):
        self.disable()

    # --------------------------------------------------------------- recording
    def _on_native(self, fn_name: str) -> None:
        if fn_name in _FUSED:
            self.fused[_FUSED[fn_name]] += 1
        if self._prev_hook is not None:
            self._prev_hook(fn_name)

    def _wrap(self, name, fn):
        log = self

        def wrapped(*args, **kwargs):
            cuda = log.sync and torch.cuda.is_available()
            if cuda:
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = fn(*args, **kwargs)
            if cuda:
                torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            nbytes = _nbytes(args, kwargs)
            group = kwargs.get("group")
            try:
                n = dist.get_world_size(group) if dist.is_initialized() else 1
            except (ValueError, RuntimeError):
                n = 1
            r = log.records[(name, nbytes)]
            r["count"] += 1
            r["bytes"] += nbytes
            r["seconds"] += dt
            r["bus_bytes"] += nbytes * _bus_factor(name, n)
            if log.verbose and (not dist.is_initialized() or dist.get_rank() == 0):
                print(f"comm op: {name} | msg size: {nbytes} B | time (ms): {dt * 1e3:.3f}", flush=True)
            return out

        wrapped.__name__ = name
        return wrapped

    # ----------------------------------------------------------------- summary
    def summary(self) -> dict:
        ops = []
        for (name, size), r in sorted(self.records.items()):
            sec = r["seconds"]
            ops.append({"op": name, "message_bytes": size, "count": r["count"], "total_ms": sec * 1e3,
                        "avg_ms": sec * 1e3 / max(1, r["count"]),
                        "algbw_gbps": r["bytes"] / sec / 1e9 if sec > 0 else 0.0,
                        "busbw_gbps": r["bus_bytes"] / sec / 1e9 if sec > 0 else 0.0})
        return {"library": ops, "fused_kernels": dict(self.fused), "timed_with_device_sync": self.sync}

    def log_summary(self, printer=print) -> None:
        s = self.summary()
        printer(f"{'comm op':<26}{'msg size (B)':>14}{'count':>8}{'total (ms)':>12}{'avg (ms)':>10}"
                f"{'algbw (GB/s)':>14}{'busbw (GB/s)':>14}")
        for o in s["library"]:
            printer(f"{o['op']:<26}{o['message_bytes']:>14}{o['count']:>8}{o['total_ms']:>12.2f}{o['avg_ms']:>10.3f}"
                    f"{o['algbw_gbps']:>14.2f}{o['busbw_gbps']:>14.2f}")
        for k, v in sorted(s["fused_kernels"].items()):
            printer(f"{k:<48}{v:>8} launches (data moved inside the kernel)")

    def reset(self) -> None:
        self.records.clear()
        self.fused.clear()
