"""Local job runner: what training-operator / mpi-operator / KubeRay / kubelet do for the
reference's charts, for one box.

A rendered `JobPlan` becomes a set of "pods" — process groups on this machine — with the
environment contract the user scripts rely on:

* static PyTorchJob : `PET_NNODES`, `PET_NPROC_PER_NODE`, `PET_NODE_RANK`, `PET_MASTER_ADDR`,
  `PET_MASTER_PORT`; pod names `<release>-master-0`, `<release>-worker-<i>`
  (charts/machine-learning/training/pytorchjob-distributed/templates/train.yaml:58-141);
* elastic PyTorchJob : `PET_RDZV_ID` (= release), `PET_RDZV_ENDPOINT`, `PET_RDZV_BACKEND`,
  `PET_NNODES=min:max` (…/pytorchjob-elastic/templates/train.yaml:58-71);
* MPIJob            : `-np` ranks with `OMPI_COMM_WORLD_{RANK,SIZE,LOCAL_RANK}` (plus the
  torch rendezvous variables so Horovod-style scripts can init NCCL);
* RayJob            : a submitter process running the script with the worker-group size and a
  local "cluster address"; `engines/ray_train_api.py` spawns the workers;
* Pod / Deployment  : one process.

Elastic jobs whose `elastic_policy` gives `min_replicas < max_replicas` keep the chart's per-pod
`restartPolicy: OnFailure` instead: a failed worker pod is restarted alone while the survivors
re-rendezvous among themselves (torchrun's c10d rendezvous admits between min and max nodes), the
launcher hosts the rendezvous store so that no pod's death takes it down, and `scale` changes the
number of worker pods inside `[min_replicas, max_replicas]` while the job runs
(charts/machine-learning/training/pytorchjob-elastic/templates/train.yaml:58-71).

GPUs are sliced per pod through `CUDA_VISIBLE_DEVICES`; `/efs`- and `/fsx`-style mount paths
are remapped under `--root`; a failed pod kills the group and the group restarts up to
`backoff_limit` times (`restartPolicy: OnFailure`); SIGINT/SIGTERM tear everything down.
State lives under `<root>/.b200flow/releases/<release>/` (plan, script, per-pod logs, status).
"""
from __future__ import annotations

import json
import os
import re
import signal
import sys
import socket
import subprocess
import time
from dataclasses import asdict
from pathlib import Path

from ..jobspec.render import JobPlan

REPO_ROOT = Path(__file__).resolve().parent.parent.parent


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def remap_paths(text: str, mounts: dict) -> str:
    """Rewrite absolute mount paths (`/efs/...`) to their local directories."""
    for src in sorted(mounts, key=len, reverse=True):
        dst = mounts[src]
        text = re.sub(r"(?<![\w./-])" + re.escape(src) + r"(?=/|\b)", dst, text)
    return text


def visible_gpu_count() -> int:
    env = os.environ.get("CUDA_VISIBLE_DEVICES")
    if env is not None:
        return len([x for x in env.split(",") if x.strip()])
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=10).stdout
        return len([l for l in out.splitlines() if l.startswith("GPU ")])
    except Exception:
        return 0


class ReleaseState:
    def __init__(self, root: Path, release: str):
        self.dir = Path(root) / ".b200flow" / "releases" / release
        self.release = release

    def exists(self) -> bool:
        return (self.dir / "plan.json").exists()

    def write_status(self, **kw) -> None:
        cur = self.read_status()
        cur.update(kw)
        cur["updated"] = time.time()
        (self.dir / "status.json").write_text(json.dumps(cur, indent=1))

    def read_status(self) -> dict:
        p = self.dir / "status.json"
        return json.loads(p.read_text()) if p.exists() else {}


class PodSpec:
    def __init__(self, name: str, argv: list[str], env: dict, gpus: list[int] | None):
        self.name, self.argv, self.env, self.gpus = name, argv, env, gpus


class Runner:
    def __init__(self, plan: JobPlan, root: str | Path, offline: bool = True, fault=None,
                 max_restarts: int | None = None, extra_env: dict | None = None):
        self.plan, self.root = plan, Path(root)
        self.state = ReleaseState(self.root, plan.release)
        self.offline = offline
        self.fault = fault
        self.max_restarts = plan.backoff_limit if max_restarts is None else max_restarts
        self.extra_env = extra_env or {}
        self.procs: list[subprocess.Popen] = []
        self.restarts = 0
        self._stop = False

    # ------------------------------------------------------------ preparation
    def prepare(self) -> Path:
        d = self.state.dir
        (d / "pods").mkdir(parents=True, exist_ok=True)
        for local in self.plan.mounts.values():
            Path(local).mkdir(parents=True, exist_ok=True)
        script = remap_paths(self.plan.script, self.plan.mounts)
        path = d / self.plan.script_name
        path.write_text(script)
        path.chmod(0o555 | 0o200)      # the chart mounts it 0555 (train.yaml:80, defaultMode 365)
        (d / "plan.json").write_text(json.dumps(asdict(self.plan), indent=1, default=str))
        if self.offline:
            from .offline import write_shims

            write_shims(d / "shims")
        self.state.write_status(phase="Pending", restarts=0, chart=self.plan.chart, kind=self.plan.kind)
        return path

    def _base_env(self, pod: str) -> dict:
        env = dict(os.environ)
        for k, v in self.plan.env.items():
            env[k] = remap_paths(v, self.plan.mounts)
        env.update({k: str(v) for k, v in self.extra_env.items()})
        env["HOSTNAME"] = pod
        env["B2F_RELEASE"] = self.plan.release
        env["B2F_ROOT"] = str(self.root)
        env["B2F_MOUNTS"] = json.dumps(self.plan.mounts)
        env["PYTHONPATH"] = str(REPO_ROOT) + os.pathsep + env.get("PYTHONPATH", "")
        env.setdefault("MASTER_ADDR", "127.0.0.1")
        if self.offline:
            env["PATH"] = str(self.state.dir / "shims") + os.pathsep + env.get("PATH", "")
            env["B2F_OFFLINE"] = "1"
            env["PYTHONPATH"] = str(REPO_ROOT / "b200flow" / "workloads" / "offline_site") + os.pathsep + \
                env["PYTHONPATH"]
        if "HOME" in env:
            Path(env["HOME"]).mkdir(parents=True, exist_ok=True)
        if self.fault:
            env["B2F_INJECT_FAULT"] = self.fault
        env["B2F_RESTART_COUNT"] = str(self.restarts)
        return env

    def _gpu_slices(self, npods: int) -> list[list[int] | None]:
        total = visible_gpu_count()
        per = self.plan.gpus_per_node
        if not total or not per:
            return [None] * npods
        base = os.environ.get("CUDA_VISIBLE_DEVICES")
        ids = [x.strip() for x in base.split(",")] if base else [str(i) for i in range(total)]
        out = []
        for i in range(npods):
            sl = ids[(i * per) % total:(i * per) % total + per]
            out.append(sl if len(sl) == per else ids[:min(per, total)])
        return out

    def spec_error(self) -> str | None:
        """A job spec this box can never run, found before anything is spawned."""
        p = self.plan
        if p.kind in ("static", "elastic") and p.gpus_per_node and p.nproc_per_node and visible_gpu_count() \
                and int(p.nproc_per_node) > int(p.gpus_per_node):
            # a pod sees only the GPUs of its `nvidia.com/gpu` limit: more ranks than that would share devices and
            # die inside NCCL ("Duplicate GPU detected") after the whole pre-script has run
            return (f"resources.nproc_per_node={p.nproc_per_node} exceeds the pod's GPU limit "
                    f"(resources.limits.\"nvidia.com/gpu\"={p.gpus_per_node}); raise the limit, e.g. "
                    f"--set 'resources.limits.nvidia\\.com/gpu={p.nproc_per_node}'")
        return None

    def pods(self) -> list[PodSpec]:
        p, script = self.plan, str(self.state.dir / self.plan.script_name)
        sh = ["bash", script]      # the data-process chart says `sh`, which is bash on its images
        rel = p.release
        specs: list[PodSpec] = []
        if p.kind == "static":
            port = free_port()
            slices = self._gpu_slices(p.nnodes)
            for r in range(p.nnodes):
                name = f"{rel}-master-0" if r == 0 else f"{rel}-worker-{r - 1}"
                env = self._base_env(name)
                env.update(PET_NNODES=str(p.nnodes), PET_NODE_RANK=str(r), PET_MASTER_ADDR="127.0.0.1",
                           PET_MASTER_PORT=str(port))
                if p.nproc_per_node:
                    env["PET_NPROC_PER_NODE"] = str(p.nproc_per_node)
                specs.append(PodSpec(name, sh, env, slices[r]))
        elif p.kind == "elastic":
            port = free_port()
            for r in range(p.nnodes):
                specs.append(self._elastic_pod(r, port, max(p.nnodes, 1)))
        elif p.kind in ("mpijob", "maskrcnn"):
            np_ = p.mpirun_np or 1
            total = visible_gpu_count()
            if total and np_ > total:
                np_ = total            # one box: cap ranks at the GPUs that exist
            port = free_port()
            per = max(1, p.gpus_per_node or 1)
            for r in range(np_):
                name = f"{rel}-worker-{r // per}"
                env = self._base_env(name)
                env.update(OMPI_COMM_WORLD_RANK=str(r), OMPI_COMM_WORLD_SIZE=str(np_),
                           OMPI_COMM_WORLD_LOCAL_RANK=str(r % per), OMPI_COMM_WORLD_LOCAL_SIZE=str(min(per, np_)),
                           RANK=str(r), WORLD_SIZE=str(np_), LOCAL_RANK=str(r % per),
                           MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                           B2F_MPIRUN_ARGS=json.dumps(p.mpirun_args))
                pod = PodSpec(f"{name}-rank{r}", sh, env, None)
                specs.append(pod)
        elif p.kind == "rayjob":
            name = f"{rel}-submitter"
            env = self._base_env(name)
            ray = p.ray or {}
            env.update(RAY_ADDRESS="local", B2F_RAY_WORKERS=str(ray.get("workers", 1)),
                       B2F_RAY_GPUS_PER_WORKER=str(p.gpus_per_node or 0),
                       B2F_RAY_RUNTIME_ENV=str(ray.get("runtime_env_yaml") or ""))
            specs.append(PodSpec(name, sh, env, None))
        else:  # pod / deployment
            name = f"{rel}-pod"
            specs.append(PodSpec(name, sh, self._base_env(name), self._gpu_slices(1)[0]))
        return specs

    def _elastic_bounds(self) -> tuple[int, int]:
        el = self.plan.elastic or {}
        lo, hi = int(el.get("min_replicas", 1)), int(el.get("max_replicas", 1))
        return lo, max(lo, hi)

    def _elastic_pod(self, r: int, port: int, npods: int) -> PodSpec:
        p = self.plan
        el = p.elastic or {}
        lo, hi = self._elastic_bounds()
        name = f"{p.release}-worker-{r}"
        env = self._base_env(name)
        env.update(PET_NNODES=f"{lo}:{hi}" if lo != hi else str(p.nnodes),
                   PET_RDZV_ID=p.release, PET_RDZV_BACKEND=str(el.get("rdzv_backend", "c10d")),
                   PET_RDZV_ENDPOINT=f"127.0.0.1:{port}")
        if p.nproc_per_node:
            env["PET_NPROC_PER_NODE"] = str(p.nproc_per_node)
        if lo != hi:
            # torch's dynamic rendezvous caches the trainers' bootstrap store of the FIRST round it joined; a node
            # that joins a later round would wait forever for an address nobody publishes again.  Without the
            # shared store every round publishes its own (rank 0 sets, the others get).
            env.setdefault("TORCH_DISABLE_SHARE_RDZV_TCP_STORE", "1")
        slices = self._gpu_slices(max(npods, r + 1))
        return PodSpec(name, ["bash", str(self.state.dir / p.script_name)], env, slices[r])

    # ------------------------------------------------------------- execution
    def _spawn(self, spec: PodSpec) -> subprocess.Popen:
        env = dict(spec.env)
        if spec.gpus is not None:
            env["CUDA_VISIBLE_DEVICES"] = ",".join(spec.gpus)
        logdir = self.state.dir / "pods" / spec.name
        logdir.mkdir(parents=True, exist_ok=True)
        log = open(logdir / "stdout.log", "ab")
        log.write(f"--- start (restart {self.restarts}) {time.strftime('%F %T')}\n".encode())
        log.flush()
        cwd = env.get("HOME") if env.get("HOME") and Path(env["HOME"]).is_dir() else None
        return subprocess.Popen(spec.argv, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=cwd,
                                start_new_session=True)

    def _kill_all(self, sig=signal.SIGTERM) -> None:
        for pr in self.procs:
            if pr.poll() is None:
                try:
                    os.killpg(pr.pid, sig)
                except ProcessLookupError:
                    pass
        deadline = time.time() + 10
        for pr in self.procs:
            try:
                pr.wait(timeout=max(0.1, deadline - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(pr.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass

    def stop(self) -> None:
        """Take the release's pods down and make `run` return (what `helm uninstall` does to a running job)."""
        self._stop = True
        self._kill_all()

    def _on_signal(self, signum, frame):
        self.stop()

    def _log_progress(self) -> int:
        """Total bytes the pods have logged so far: the job's heartbeat."""
        total = 0
        for f in (self.state.dir / "pods").glob("*/stdout.log"):
            try:
                total += f.stat().st_size
            except OSError:
                pass
        return total

    # ------------------------------------------------------------- admission
    def gpus_needed(self) -> int:
        p = self.plan
        per = int(p.gpus_per_node or 0)
        if p.kind in ("mpijob", "maskrcnn"):
            return min(int(p.mpirun_np or 1), visible_gpu_count()) if per else 0
        if p.kind == "elastic":
            return per * min(self._elastic_bounds()[1], max(self._elastic_bounds()[0], int(p.nnodes or 1)))
        if p.kind == "rayjob":
            return per * int((p.ray or {}).get("workers", 1))
        return per * int(p.nnodes or 1) if p.kind == "static" else per

    def _gpus_claimed_by_others(self) -> int:
        total = 0
        for f in (self.root / ".b200flow" / "releases").glob("*/status.json"):
            if f.parent.name == self.plan.release:
                continue
            try:
                st = json.loads(f.read_text())
                if st.get("phase") in ("Running", "Restarting") and st.get("runner_pid"):
                    os.kill(int(st["runner_pid"]), 0)          # a dead runner holds nothing
                    total += int(st.get("gpus_claimed", 0))
            except (OSError, ValueError, ProcessLookupError, PermissionError):
                continue
        return total

    def _admit(self, pending_timeout: float | None, poll: float) -> bool:
        """A release stays Pending while the GPUs its pods request are held by other running releases under the
        same root (the scheduler's side of a pod's `Pending` phase); over-subscribed requests on a box with fewer
        GPUs than asked are admitted when nothing else runs, as before."""
        need, have = self.gpus_needed(), visible_gpu_count()
        t0 = time.time()
        while need and have and not self._stop:
            used = self._gpus_claimed_by_others()
            if used == 0 or used + need <= have:
                break
            if pending_timeout is not None and time.time() - t0 > pending_timeout:
                return False
            time.sleep(poll)
        self.state.write_status(gpus_claimed=need if have else 0, runner_pid=os.getpid())
        return not self._stop

    def run(self, timeout: float | None = None, poll: float = 0.2, hang_timeout: float | None = None,
            pending_timeout: float | None = None) -> int:
        """Run to completion. Returns 0 on success, the failing exit code otherwise.

        `pending_timeout`: seconds the release may wait for GPUs held by other releases before it fails (exit 126).

        `hang_timeout`: seconds without a single new log byte from any pod after which the group is declared
        hung (exit code 125), killed and restarted like any other failure — the local stand-in for the NCCL
        watchdog / lost-heartbeat detection a cluster would provide."""
        self.prepare()
        err = self.spec_error()
        if err:
            print(f"[b200flow] {self.plan.release}: {err}", file=sys.stderr, flush=True)
            self.state.write_status(phase="Failed", exit_code=2, message=err)
            return 2
        old = {}
        for s in (signal.SIGINT, signal.SIGTERM):
            try:
                old[s] = signal.signal(s, self._on_signal)
            except ValueError:      # not the main thread
                pass
        if not self._admit(pending_timeout, max(poll, 0.2)):
            self.state.write_status(phase="Failed", exit_code=126, reason="Pending: GPUs not available in time")
            for s, h in old.items():
                signal.signal(s, h)
            return 126
        t0 = time.time()
        rc = 1
        lo, hi = self._elastic_bounds()
        if self.plan.kind == "elastic" and lo < hi:
            try:
                return self._run_elastic(lo, hi, timeout, poll, t0)
            finally:
                self._kill_all()
                for s, h in old.items():
                    signal.signal(s, h)
        try:
            while not self._stop:
                specs = self.pods()
                self.procs = [self._spawn(s) for s in specs]
                self.state.write_status(phase="Running", restarts=self.restarts,
                                        pods={s.name: p.pid for s, p in zip(specs, self.procs)})
                failed = None
                seen, seen_at = self._log_progress(), time.time()
                while True:
                    codes = [p.poll() for p in self.procs]
                    bad = [c for c in codes if c not in (None, 0)]
                    if bad:
                        failed = bad[0]
                        break
                    if all(c == 0 for c in codes):
                        break
                    if timeout is not None and time.time() - t0 > timeout:
                        failed = 124
                        self._stop = True
                        break
                    if hang_timeout is not None:
                        now_bytes = self._log_progress()
                        if now_bytes != seen:
                            seen, seen_at = now_bytes, time.time()
                        elif time.time() - seen_at > hang_timeout:
                            failed = 125
                            break
                    if self._stop:
                        failed = 130
                        break
                    time.sleep(poll)
                if failed is None:
                    rc = 0
                    self.state.write_status(phase="Succeeded", exit_code=0)
                    break
                self._kill_all()                      # cleanPodPolicy: take the whole group down
                rc = failed
                if self._stop or self.restarts >= self.max_restarts:
                    self.state.write_status(phase="Failed", exit_code=failed, restarts=self.restarts)
                    break
                self.restarts += 1
                self.state.write_status(phase="Restarting", restarts=self.restarts, last_exit=failed)
                time.sleep(min(10.0, 0.05 * (2 ** min(self.restarts, 8))))   # crash-loop back-off
        finally:
            self._kill_all()
            for s, h in old.items():
                signal.signal(s, h)
        return rc


    # ----------------------------------------- elastic jobs with membership change (min_replicas < max_replicas)
    def _start_rdzv_store(self, port: int):
        """The c10d rendezvous store lives with the launcher (the operator's side), not inside a worker pod, so any
        pod may die.  torchrun's c10d backend falls back to a client connection when the port is already served."""
        code = ("import sys, time\nfrom datetime import timedelta\nfrom torch.distributed import TCPStore\n"
                f"s = TCPStore('127.0.0.1', {port}, None, True, timedelta(seconds=30), wait_for_workers=False)\n"
                "print('rdzv store up', flush=True)\n"
                "while True:\n    time.sleep(3600)\n")
        log = open(self.state.dir / "rdzv_store.log", "ab")
        pr = subprocess.Popen([os.environ.get("B2F_PYTHON", "python"), "-c", code], stdout=log,
                              stderr=subprocess.STDOUT, start_new_session=True)
        deadline = time.time() + 120
        while time.time() < deadline and pr.poll() is None:
            try:
                with socket.create_connection(("127.0.0.1", port), timeout=0.5):
                    return pr
            except OSError:
                time.sleep(0.1)
        return pr

    def scale_request(self) -> int | None:
        f = self.state.dir / "scale"
        try:
            return int(f.read_text().strip())
        except (OSError, ValueError):
            return None

    def _run_elastic(self, lo: int, hi: int, timeout, poll, t0) -> int:
        p = self.plan
        port = free_port()
        backend = str((p.elastic or {}).get("rdzv_backend", "c10d"))
        store = self._start_rdzv_store(port) if backend == "c10d" else None
        desired = min(hi, max(lo, p.nnodes))
        pods: dict[int, subprocess.Popen] = {}        # worker index -> running process
        respawn_at: dict[int, float] = {}             # failed workers waiting out their back-off
        done: set[int] = set()                        # exited 0
        removed: set[int] = set()                     # scaled away on request
        pod_restarts: dict[int, int] = {}
        changes = 0
        (self.state.dir / "scale").unlink(missing_ok=True)

        def spawn(r: int):
            spec = self._elastic_pod(r, port, hi)
            spec.env["B2F_RESTART_COUNT"] = str(pod_restarts.get(r, 0))
            pods[r] = self._spawn(spec)
            self.procs = list(pods.values()) + ([store] if store is not None else [])

        def publish(phase="Running"):
            self.state.write_status(phase=phase, restarts=self.restarts, replicas=len(pods) + len(respawn_at),
                                    desired_replicas=desired, min_replicas=lo, max_replicas=hi,
                                    membership_changes=changes,
                                    pods={f"{p.release}-worker-{r}": pr.pid for r, pr in pods.items()})

        for r in range(desired):
            spawn(r)
        publish()
        rc = 1
        try:
            while True:
                if self._stop:
                    rc = 130
                    break
                if timeout is not None and time.time() - t0 > timeout:
                    rc, self._stop = 124, True
                    break
                want = self.scale_request()
                if want is not None and min(hi, max(lo, want)) != desired:
                    desired = min(hi, max(lo, want))
                live = sorted(set(pods) | set(respawn_at))
                if len(live) + len(done) > desired and live and not done:        # scale in: highest indices leave
                    r = live[-1]
                    removed.add(r)
                    respawn_at.pop(r, None)
                    pr = pods.pop(r, None)
                    if pr is not None and pr.poll() is None:
                        try:
                            os.killpg(pr.pid, signal.SIGTERM)
                        except ProcessLookupError:
                            pass
                    changes += 1
                    publish()
                elif len(live) < desired and not done:                            # scale out: lowest free index joins
                    r = next(i for i in range(hi + 1) if i not in pods and i not in respawn_at)
                    removed.discard(r)
                    spawn(r)
                    changes += 1
                    publish()
                for r, pr in list(pods.items()):
                    code = pr.poll()
                    if code is None:
                        continue
                    del pods[r]
                    if code == 0:
                        done.add(r)
                        continue
                    self.restarts += 1
                    pod_restarts[r] = pod_restarts.get(r, 0) + 1
                    changes += 1
                    if self.restarts > self.max_restarts:
                        rc = code
                        self.state.write_status(phase="Failed", exit_code=code, restarts=self.restarts)
                        return rc
                    # restartPolicy OnFailure: this pod alone comes back; the others re-rendezvous meanwhile
                    respawn_at[r] = time.time() + min(10.0, 0.5 * (2 ** min(pod_restarts[r], 4)))
                    publish()
                for r, at in list(respawn_at.items()):
                    if time.time() >= at:
                        del respawn_at[r]
                        spawn(r)
                        publish()
                if not pods and not respawn_at:
                    rc = 0 if done else 1
                    break
                time.sleep(poll)
            self.state.write_status(phase="Succeeded" if rc == 0 else "Failed", exit_code=rc, restarts=self.restarts,
                                    membership_changes=changes)
            return rc
        finally:
            self.procs = list(pods.values()) + ([store] if store is not None else [])


def scale(root: str | Path, release: str, replicas: int) -> bool:
    """Ask a running elastic release for `replicas` worker pods (clamped to its min..max by the runner)."""
    st = ReleaseState(Path(root), release)
    if not st.exists():
        return False
    tmp = st.dir / ".scale.tmp"
    tmp.write_text(str(int(replicas)))
    os.replace(tmp, st.dir / "scale")
    return True


def uninstall(root: str | Path, release: str) -> bool:
    """Kill whatever is left of a release and mark it uninstalled (logs are kept)."""
    st = ReleaseState(Path(root), release)
    if not st.exists():
        return False
    cur = st.read_status()
    sup = cur.get("supervisor_pid")
    if sup:        # a release installed with --no-wait: its supervisor takes the pods down when it is told to stop
        try:
            os.kill(int(sup), signal.SIGTERM)
            for _ in range(100):
                os.kill(int(sup), 0)
                time.sleep(0.1)
        except (ProcessLookupError, PermissionError):
            pass
    for name, pid in (st.read_status().get("pods") or {}).items():
        try:
            os.killpg(int(pid), signal.SIGTERM)
        except (ProcessLookupError, PermissionError):
            pass
    st.write_status(phase="Uninstalled", supervisor_pid=None)
    return True


def status(root: str | Path, release: str) -> dict:
    return ReleaseState(Path(root), release).read_status()


def logs(root: str | Path, release: str, tail: int | None = None) -> dict:
    st = ReleaseState(Path(root), release)
    out = {}
    for f in sorted((st.dir / "pods").glob("*/stdout.log")):
        txt = f.read_text(errors="replace")
        out[f.parent.name] = "\n".join(txt.splitlines()[-tail:]) if tail else txt
    return out
