"""Metrics / logging: JSONL + TensorBoard scalars, rank-aware text logs, CUDA-event timers,
fault injection.

The reference gets TensorBoard event files from Accelerate's `--with_tracking --project_dir`
(examples/accelerate/bert-glue-mrpc/pretrain.yaml:48,50) and from tensorpack under
`…/train_log/maskrcnn` (served by the *-jupyter chart), and a Megatron log line every
`--log-interval`.  `Tracker` writes both a `metrics.jsonl` and — when
the `tensorboard` package is importable — event files under `<project_dir>/<run_name>`.
"""
from __future__ import annotations

import json
import os
import signal
import sys
import time
from pathlib import Path

import torch


class Tracker:
    def __init__(self, project_dir: str | os.PathLike, run_name: str = "run", config: dict | None = None):
        self.dir = Path(project_dir) / run_name
        self.dir.mkdir(parents=True, exist_ok=True)
        self.jsonl = open(self.dir / "metrics.jsonl", "a")
        self.tb = None
        try:
            from torch.utils.tensorboard import SummaryWriter

            self.tb = SummaryWriter(log_dir=str(self.dir))
        except Exception:
            self.tb = None
        if config:
            (self.dir / "hparams.json").write_text(json.dumps(config, default=str, indent=1))

    def log(self, values: dict, step: int | None = None) -> None:
        rec = {"step": step, "time": time.time(), **{k: _num(v) for k, v in values.items()}}
        self.jsonl.write(json.dumps(rec) + "\n")
        self.jsonl.flush()
        if self.tb is not None:
            for k, v in values.items():
                v = _num(v)
                if isinstance(v, (int, float)):
                    self.tb.add_scalar(k, v, global_step=step)

    def finish(self) -> None:
        self.jsonl.close()
        if self.tb is not None:
            self.tb.flush()
            self.tb.close()


def _num(v):
    if isinstance(v, torch.Tensor):
        return v.item() if v.numel() == 1 else v.tolist()
    return v


class Timeline:
    """Chrome-trace event log (`chrome://tracing` / Perfetto), the `HOROVOD_TIMELINE=<file>` facility of the
    reference's Horovod stack: host-side instants and spans, one process row per rank."""

    def __init__(self, path: str | os.PathLike, rank: int = 0):
        self.path, self.rank, self.events, self._t0 = Path(path), rank, [], time.perf_counter()

    def _us(self) -> float:
        return (time.perf_counter() - self._t0) * 1e6

    def instant(self, name: str, **args) -> None:
        self.events.append({"name": name, "ph": "i", "s": "t", "ts": self._us(), "pid": self.rank, "tid": 0,
                            "args": args})

    def span(self, name: str, **args):
        tl = self

        class _Span:
            def __enter__(self):
                self.t = tl._us()

            def __exit__(self, *exc):
                tl.events.append({"name": name, "ph": "X", "ts": self.t, "dur": tl._us() - self.t,
                                  "pid": tl.rank, "tid": 0, "args": args})
        return _Span()

    def flush(self) -> None:
        self.path.parent.mkdir(parents=True, exist_ok=True)
        self.path.write_text(json.dumps({"traceEvents": self.events, "displayTimeUnit": "ms"}))


class Timers:
    """Named CUDA-event timers (wall-clock on CPU); `summary(group)` reduces max over ranks —
    the Megatron `Timers` / DeepSpeed `wall_clock_breakdown` equivalent."""

    def __init__(self, use_cuda: bool | None = None):
        self.use_cuda = torch.cuda.is_available() if use_cuda is None else use_cuda
        self._open, self._acc, self._count = {}, {}, {}

    def start(self, name: str) -> None:
        if self.use_cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._open[name] = e
        else:
            self._open[name] = time.perf_counter()

    def stop(self, name: str) -> None:
        s = self._open.pop(name)
        if self.use_cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._acc.setdefault(name, []).append((s, e))
        else:
            self._acc[name] = self._acc.get(name, 0.0) + (time.perf_counter() - s) * 1e3
        self._count[name] = self._count.get(name, 0) + 1

    def elapsed_ms(self, reset: bool = True) -> dict:
        out = {}
        if self.use_cuda:
            torch.cuda.synchronize()
            for k, pairs in self._acc.items():
                out[k] = sum(s.elapsed_time(e) for s, e in pairs)
        else:
            out = dict(self._acc)
        if reset:
            self._acc, self._count = {}, {}
        return out

    def summary(self, group=None, reset: bool = True) -> dict:
        import torch.distributed as dist

        ms = self.elapsed_ms(reset)
        if dist.is_available() and dist.is_initialized() and ms:
            keys = sorted(ms)
            dev = "cuda" if (self.use_cuda and dist.get_backend(group) == "nccl") else "cpu"
            t = torch.tensor([ms[k] for k in keys], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            ms = dict(zip(keys, t.tolist()))
        return ms


def rank_log_path(home: str | os.PathLike, node: str, name: str = "train.log") -> Path:
    """`$HOME/logs/<node>/<name>` — the per-pod log convention of the examples
    (examples/accelerate/bert-glue-mrpc/pretrain.yaml:28-30)."""
    p = Path(home) / "logs" / str(node)
    p.mkdir(parents=True, exist_ok=True)
    return p / name


def maybe_inject_fault(step: int, rank: int | None = None) -> None:
    """`--inject-fault rank=R,step=S,kind={exit,hang,sigkill}` (launcher flag -> env var).
    Fires only on the first incarnation of the job so restart tests terminate."""
    spec = os.environ.get("B2F_INJECT_FAULT")
    if not spec or os.environ.get("B2F_RESTART_COUNT", "0") != "0":
        return
    kv = dict(x.split("=") for x in spec.split(","))
    r = int(os.environ.get("RANK", "0")) if rank is None else rank
    if int(kv.get("rank", 0)) != r or int(kv.get("step", -1)) != step:
        return
    kind = kv.get("kind", "exit")
    print(f"[b200flow] injected fault {kind} at rank {r} step {step}", flush=True)
    if kind == "exit":
        sys.exit(17)
    if kind == "sigkill":
        os.kill(os.getpid(), signal.SIGKILL)
    if kind == "hang":
        while True:
            time.sleep(3600)
