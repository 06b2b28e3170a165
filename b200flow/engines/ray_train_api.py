"""Ray-Train-shaped API (`TorchTrainer`, `ScalingConfig`, `RunConfig`, `CheckpointConfig`,
`Checkpoint`, `report`, `get_dataset_shard`, `get_context`) on a local worker group.

The reference submits `fine-tune-pytorch-text-classifier.py` to a KubeRay cluster
(charts/machine-learning/training/raytrain/templates/train.yaml:49-62;
examples/ray/lightning-bert/fine-tune.yaml:46-49; call stack SURVEY.md §3.3).  What that
workload exercises of Ray is: a worker group with a torch process group, dataset shards
handed to the workers, and `train.report(metrics, checkpoint)` flowing back to a driver
that persists checkpoints under `storage_path` keeping the best `num_to_keep`.

Here the driver forks `num_workers` processes on the box (one GPU each when `use_gpu`),
hands each its dataset shard and receives reports through the C++ shared-memory mailbox
(`csrc/host/shm_channel.cpp`) — the native stand-in for the object-store hand-off.
"""
from __future__ import annotations

import ctypes
import multiprocessing as mp
import os
import pickle
import shutil
import socket
import time
import traceback
from dataclasses import dataclass, field
from pathlib import Path


# ------------------------------------------------------------------ mailbox
class Channel:
    """SPSC byte channel over POSIX shared memory (libb200flow_host.so)."""

    def __init__(self, name: str, create: bool, capacity: int = 64 << 20):
        from ..ops._lib import host_lib

        self.lib = host_lib()
        self.lib.b2f_chan_create.restype = ctypes.c_void_p
        self.lib.b2f_chan_open.restype = ctypes.c_void_p
        self.lib.b2f_chan_recv.restype = ctypes.c_longlong
        self.name = name.encode()
        if create:
            self.h = self.lib.b2f_chan_create(self.name, ctypes.c_uint64(capacity))
        else:
            self.h = None
            for _ in range(200):
                self.h = self.lib.b2f_chan_open(self.name)
                if self.h:
                    break
                time.sleep(0.05)
        if not self.h:
            raise OSError(f"cannot {'create' if create else 'open'} shared-memory channel {name}")
        self._buf = ctypes.create_string_buffer(1 << 20)

    def send(self, obj, timeout_ms: int = 60000) -> None:
        data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        rc = self.lib.b2f_chan_send(ctypes.c_void_p(self.h), data, ctypes.c_uint32(len(data)),
                                    ctypes.c_long(timeout_ms))
        if rc != 0:
            raise OSError(f"channel send failed rc={rc} ({len(data)} bytes)")

    def recv(self, timeout_ms: int = 100):
        need = ctypes.c_uint32(0)
        n = self.lib.b2f_chan_recv(ctypes.c_void_p(self.h), self._buf, ctypes.c_uint32(len(self._buf)),
                                   ctypes.c_long(timeout_ms), ctypes.byref(need))
        if n == -4:
            self._buf = ctypes.create_string_buffer(int(need.value) + 1024)
            return self.recv(timeout_ms)
        if n == -2:
            return None
        if n < 0:
            raise EOFError(f"channel closed rc={n}")
        return pickle.loads(self._buf.raw[:n])

    def close(self):
        if self.h:
            self.lib.b2f_chan_close(ctypes.c_void_p(self.h))

    def free(self):
        if self.h:
            self.lib.b2f_chan_free(ctypes.c_void_p(self.h))
            self.h = None


# ------------------------------------------------------------------- config
@dataclass
class ScalingConfig:
    num_workers: int = 1
    use_gpu: bool = False
    resources_per_worker: dict | None = None


@dataclass
class CheckpointConfig:
    num_to_keep: int | None = None
    checkpoint_score_attribute: str | None = None
    checkpoint_score_order: str = "max"


@dataclass
class FailureConfig:
    """`max_failures`: how often a failed worker group is started again (from the newest reported checkpoint,
    which the training function picks up with `get_checkpoint()`); -1 = without limit."""
    max_failures: int = 0


@dataclass
class RunConfig:
    name: str | None = None
    storage_path: str | None = None
    checkpoint_config: CheckpointConfig = field(default_factory=CheckpointConfig)
    failure_config: FailureConfig = field(default_factory=FailureConfig)


class Checkpoint:
    def __init__(self, path: str):
        self.path = str(path)

    @classmethod
    def from_directory(cls, path: str) -> "Checkpoint":
        return cls(path)

    def as_directory(self):
        import contextlib

        @contextlib.contextmanager
        def cm():
            yield self.path
        return cm()

    def to_directory(self, dst: str) -> str:
        shutil.copytree(self.path, dst, dirs_exist_ok=True)
        return dst


@dataclass
class Result:
    metrics: dict | None
    checkpoint: Checkpoint | None
    path: str
    error: str | None = None
    best_checkpoints: list = field(default_factory=list)
    metrics_history: list = field(default_factory=list)

    @property
    def metrics_dataframe(self):
        import pandas as pd

        return pd.DataFrame(self.metrics_history)


class ListDataset:
    """Minimal `ray.data.Dataset`: rows, `map_batches`, `split`, `iter_torch_batches`."""

    def __init__(self, rows: list[dict]):
        self.rows = rows

    @classmethod
    def from_items(cls, rows):
        return cls(list(rows))

    def count(self) -> int:
        return len(self.rows)

    def map_batches(self, fn, batch_size: int = 1024, **kw):
        out = []
        for i in range(0, len(self.rows), batch_size):
            chunk = self.rows[i:i + batch_size]
            cols = {k: [r[k] for r in chunk] for k in chunk[0]}
            res = fn(cols)
            n = len(next(iter(res.values())))
            out += [{k: res[k][j] for k in res} for j in range(n)]
        return ListDataset(out)

    def split(self, n: int) -> list["ListDataset"]:
        return [ListDataset(self.rows[i::n]) for i in range(n)]

    def iter_torch_batches(self, batch_size: int = 32, collate_fn=None, drop_last: bool = False):
        import torch

        for i in range(0, len(self.rows), batch_size):
            chunk = self.rows[i:i + batch_size]
            if drop_last and len(chunk) < batch_size:
                break
            if collate_fn is not None:
                yield collate_fn(chunk)
            else:
                yield {k: torch.as_tensor([r[k] for r in chunk]) for k in chunk[0]}


# ---------------------------------------------------------- worker context
class _Context:
    def __init__(self, rank, world, local_rank, shards, report_chan, run_dir, checkpoint=None):
        self.rank, self.world, self.local_rank = rank, world, local_rank
        self.shards, self.report_chan, self.run_dir = shards, report_chan, run_dir
        self.iteration = 0
        self.checkpoint = checkpoint

    def get_world_rank(self):
        return self.rank

    def get_world_size(self):
        return self.world

    def get_local_rank(self):
        return self.local_rank

    def get_trial_dir(self):
        return self.run_dir


_ctx: _Context | None = None


def get_context() -> _Context:
    assert _ctx is not None, "not inside a TorchTrainer worker"
    return _ctx


def get_checkpoint() -> Checkpoint | None:
    """The checkpoint to continue from: `TorchTrainer(resume_from_checkpoint=...)`, or the newest one reported
    before a worker failure when the run was restarted under `FailureConfig`."""
    return get_context().checkpoint


def get_dataset_shard(name: str) -> ListDataset:
    return get_context().shards[name]


def report(metrics: dict, checkpoint: Checkpoint | None = None) -> None:
    c = get_context()
    c.iteration += 1
    ck = None
    if checkpoint is not None and c.rank == 0:
        ck = checkpoint.path
    c.report_chan.send({"rank": c.rank, "iteration": c.iteration, "metrics": dict(metrics), "checkpoint": ck})
    if checkpoint is not None:
        # the driver copies the directory before the worker may delete/reuse it
        import torch.distributed as dist

        if dist.is_initialized():
            dist.barrier()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker_main(rank, world, use_gpu, port, train_fn, config, run_id, run_dir):
    global _ctx
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    rc = 0
    chan = None
    try:
        import torch
        import torch.distributed as dist

        if use_gpu and torch.cuda.is_available():
            torch.cuda.set_device(rank % torch.cuda.device_count())
        dist.init_process_group("nccl" if (use_gpu and torch.cuda.is_available()) else "gloo",
                                rank=rank, world_size=world)
        data_chan = Channel(f"/b2f_{run_id}_data_{rank}", create=False)
        shards = None
        while shards is None:
            shards = data_chan.recv(1000)
        chan = Channel(f"/b2f_{run_id}_report_{rank}", create=False)
        start = shards.pop("__checkpoint__", None)
        _ctx = _Context(rank, world, rank, shards, chan, run_dir, Checkpoint(start) if start else None)
        train_fn(config) if config is not None else train_fn()
        chan.send({"rank": rank, "done": True})
        dist.barrier()
        dist.destroy_process_group()
    except BaseException as e:  # noqa: BLE001
        rc = 1
        msg = "".join(traceback.format_exception(type(e), e, e.__traceback__))
        try:
            (chan or Channel(f"/b2f_{run_id}_report_{rank}", create=False)).send(
                {"rank": rank, "error": msg})
        except Exception:
            print(msg, flush=True)
    os._exit(rc)


class TorchTrainer:
    def __init__(self, train_loop_per_worker, *, train_loop_config: dict | None = None,
                 scaling_config: ScalingConfig | None = None, run_config: RunConfig | None = None,
                 datasets: dict | None = None, resume_from_checkpoint: Checkpoint | None = None):
        self.resume = resume_from_checkpoint
        self.fn, self.config = train_loop_per_worker, train_loop_config
        self.scaling = scaling_config or ScalingConfig()
        self.run = run_config or RunConfig()
        self.datasets = datasets or {}

    def fit(self) -> Result:
        """Run the worker group; on a worker failure start it again (at most `failure_config.max_failures` times)
        from the newest checkpoint reported so far."""
        budget = self.run.failure_config.max_failures
        start = self.resume.path if self.resume is not None else None
        history: list = []
        attempt = 0
        while True:
            try:
                res = self._fit_once(start, attempt)
                res.metrics_history = history + res.metrics_history
                return res
            except _GroupFailed as e:
                history += e.history
                if budget >= 0 and attempt >= budget:
                    raise RuntimeError(f"TorchTrainer worker failed:\n{e.error}") from None
                attempt += 1
                start = e.last_checkpoint or start
                print(f"[b200flow] worker group failed; restart {attempt} from {start or 'scratch'}", flush=True)

    def _fit_once(self, start_checkpoint: str | None, attempt: int) -> Result:
        n = int(os.environ.get("B2F_RAY_WORKERS_OVERRIDE", self.scaling.num_workers))
        name = self.run.name or f"TorchTrainer_{time.strftime('%Y-%m-%d_%H-%M-%S')}"
        root = Path(self.run.storage_path or os.path.expanduser("~/ray_results")) / name
        root.mkdir(parents=True, exist_ok=True)
        run_id = f"{os.getpid()}_{int(time.time() * 1000) % 100000}"
        port = _free_port()
        data_ch = [Channel(f"/b2f_{run_id}_data_{r}", create=True, capacity=256 << 20) for r in range(n)]
        rep_ch = [Channel(f"/b2f_{run_id}_report_{r}", create=True) for r in range(n)]
        ctx = mp.get_context("fork")
        procs = [ctx.Process(target=_worker_main,
                             args=(r, n, self.scaling.use_gpu, port, self.fn, self.config, run_id, str(root)))
                 for r in range(n)]
        for p in procs:
            p.start()
        # hand every worker its shard of every dataset ("train" is split, the rest replicated-split too)
        split = {k: (d.split(n) if hasattr(d, "split") else [d] * n) for k, d in self.datasets.items()}
        for r in range(n):
            data_ch[r].send({**{k: split[k][r] for k in split}, "__checkpoint__": start_checkpoint})
        cc = self.run.checkpoint_config
        kept: list[tuple[float, str, dict]] = []
        history, last_metrics, last_ckpt, error = [], None, None, None
        done = set()
        idx = len(list(root.glob("checkpoint_*")))       # a restarted group keeps numbering where the last one stopped
        while len(done) < n and error is None:
            progressed = False
            for r in range(n):
                if r in done:
                    continue
                try:
                    msg = rep_ch[r].recv(50)
                except EOFError:
                    msg = None
                if msg is None:
                    if not procs[r].is_alive() and r not in done:
                        error = f"worker {r} died (exit {procs[r].exitcode})"
                        break
                    continue
                progressed = True
                if msg.get("error"):
                    error = msg["error"]
                    break
                if msg.get("done"):
                    done.add(r)
                    continue
                if msg["rank"] == 0:
                    last_metrics = msg["metrics"]
                    history.append(msg["metrics"])
                    if msg.get("checkpoint"):
                        dst = root / f"checkpoint_{idx:06d}"
                        shutil.copytree(msg["checkpoint"], dst, dirs_exist_ok=True)
                        idx += 1
                        last_ckpt = Checkpoint(str(dst))
                        score = msg["metrics"].get(cc.checkpoint_score_attribute, idx) \
                            if cc.checkpoint_score_attribute else idx
                        kept.append((float(score), str(dst), msg["metrics"]))
                        if cc.num_to_keep and len(kept) > cc.num_to_keep:
                            kept.sort(key=lambda t: t[0], reverse=cc.checkpoint_score_order == "max")
                            for _, path, _m in kept[cc.num_to_keep:]:
                                shutil.rmtree(path, ignore_errors=True)
                            kept = kept[:cc.num_to_keep]
            if not progressed:
                time.sleep(0.02)
        for p in procs:
            p.join(timeout=30 if error is None else 2)
            if p.is_alive():
                p.terminate()
        for c in data_ch + rep_ch:
            c.free()
        with open(root / "result.json", "a" if attempt else "w") as f:
            f.write("".join(__import__("json").dumps(m) + "\n" for m in history))
        if last_ckpt is not None and not os.path.exists(last_ckpt.path) and kept:
            last_ckpt = Checkpoint(kept[0][1])
        res = Result(last_metrics, last_ckpt, str(root), error,
                     [(Checkpoint(p), m) for _, p, m in kept], history)
        if error:
            raise _GroupFailed(error, last_ckpt.path if last_ckpt is not None else None, history)
        return res


class _GroupFailed(Exception):
    def __init__(self, error, last_checkpoint, history):
        super().__init__(error)
        self.error, self.last_checkpoint, self.history = error, last_checkpoint, history
