"""Multi-process sample prefetching over the shared-memory mailbox.

tensorpack feeds the reference's Mask R-CNN trainers from several pre-processing processes connected by
ZMQ pipes (`MultiProcessMapData`, SURVEY.md §3.4 "get_train_dataflow ... in N procs via ZMQ"); torch's
DataLoader workers play the same role for the other workloads.  Here every worker process owns one
single-producer / single-consumer ring in POSIX shared memory (`csrc/host/shm_channel.cpp`, the same
native channel the Ray-style worker group uses) and produces the samples at positions w, w + W, w + 2W, ...
of a given index order; the consumer reads the rings round-robin, so samples arrive in exactly that order
whatever the workers' relative speed.  A ring holds `capacity` bytes, which bounds how far a worker
runs ahead.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import traceback

_STOP = "__b2f_prefetch_stop__"


def _worker(dataset, order, w, nworkers, chan_name):
    from ..engines.ray_train_api import Channel

    ch = Channel(chan_name, create=False)
    try:
        for pos in range(w, len(order), nworkers):
            ch.send(("ok", dataset[order[pos]]), timeout_ms=600000)
        ch.send((_STOP, None), timeout_ms=600000)
    except Exception:       # noqa: BLE001 - shipped to the consumer, which raises it
        try:
            ch.send(("error", traceback.format_exc()), timeout_ms=10000)
        except OSError:
            pass
    finally:
        ch.close()
        ch.free()


class PrefetchLoader:
    """Iterate `dataset[i] for i in order` with `workers` processes producing ahead.

    `workers=0` degrades to an in-process loop (tests, tiny runs).  The dataset must be picklable
    (workers are spawned, which is also safe after CUDA has been initialised in the parent)."""

    def __init__(self, dataset, order, workers: int = 2, capacity: int = 128 << 20):
        self.dataset, self.order, self.workers, self.capacity = dataset, list(order), int(workers), capacity
        self._procs, self._chans = [], []

    def __len__(self) -> int:
        return len(self.order)

    def _start(self):
        from ..engines.ray_train_api import Channel

        ctx = mp.get_context("spawn")
        tag = f"/b2f_pf_{os.getpid()}_{id(self) & 0xffffff:x}"
        for w in range(self.workers):
            name = f"{tag}_{w}"
            self._chans.append(Channel(name, create=True, capacity=self.capacity))
            p = ctx.Process(target=_worker, args=(self.dataset, self.order, w, self.workers, name), daemon=True)
            p.start()
            self._procs.append(p)

    def __iter__(self):
        if self.workers <= 0:
            for i in self.order:
                yield self.dataset[i]
            return
        self._start()
        try:
            for pos in range(len(self.order)):
                w = pos % self.workers
                msg = None
                while msg is None:
                    msg = self._chans[w].recv(timeout_ms=200)
                    if msg is None and not self._procs[w].is_alive():
                        msg = self._chans[w].recv(timeout_ms=200)      # drain what it wrote before exiting
                        if msg is None:
                            raise RuntimeError(f"prefetch worker {w} died (exit code {self._procs[w].exitcode})")
                kind, payload = msg
                if kind == "error":
                    raise RuntimeError("prefetch worker failed:\n" + payload)
                if kind == _STOP:
                    raise RuntimeError("prefetch worker ran out of samples early")
                yield payload
        finally:
            self.close()

    def close(self) -> None:
        for p in self._procs:
            if p.is_alive():
                p.terminate()
            p.join(timeout=5)
        for c in self._chans:
            c.free()
        self._procs, self._chans = [], []
