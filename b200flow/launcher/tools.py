"""Cluster utilities of the reference, for the one-box launcher.

* `smoke`      — `eks-cluster/tests/test-gpu.yaml:9-33` and `test-gpu-efa.yaml:1-43` schedule a pod that
                 asks for 8 GPUs (and EFA NICs), mount `/fsx`, and sleep so that a person can `kubectl exec`
                 in and try `nvidia-smi` / nccl-tests by hand.  Here the same questions are answered by a
                 command: are the GPUs visible, are the shared mounts there and writable, do the native
                 libraries load, does a collective over all requested ranks return the right sum (NCCL on
                 GPUs with peer access reported per pair, gloo on CPU).
* `stage-data` — `eks-cluster/utils/stage-data.yaml:1-46` copies `s3://$S3_BUCKET/$S3_PREFIX` to a directory
                 on the shared volume.  There is no object store on the box: the source is a local directory
                 (an `s3://bucket/prefix` source resolves under `$B2F_S3_ROOT/bucket/prefix`).
* `attach`     — `eks-cluster/utils/attach-pvc.yaml:1-24` starts a pod with `/efs` and `/fsx` mounted to look
                 at outputs; here it prints where those mounts live under the launcher root and what a
                 release left in them.
"""
from __future__ import annotations

import json
import os
import shutil
import subprocess
import sys
from pathlib import Path

MOUNTS = ("efs", "fsx")


def _rank_main() -> int:
    """One rank of the collective check (spawned by `smoke`)."""
    import torch
    import torch.distributed as dist

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    use_cuda = torch.cuda.is_available() and torch.cuda.device_count() >= world
    if use_cuda:
        torch.cuda.set_device(rank)
    dist.init_process_group("nccl" if use_cuda else "gloo", rank=rank, world_size=world)
    dev = torch.device("cuda", rank) if use_cuda else torch.device("cpu")
    t = torch.full((1 << 20,), float(rank + 1), device=dev)
    dist.all_reduce(t)
    ok = bool((t == world * (world + 1) / 2).all())
    peers = {}
    if use_cuda:
        peers = {str(p): bool(torch.cuda.can_device_access_peer(rank, p)) for p in range(world) if p != rank}
    out = [None] * world
    dist.all_gather_object(out, {"rank": rank, "allreduce_ok": ok, "peer_access": peers})
    if rank == 0:
        print("B2F_SMOKE " + json.dumps({"backend": "nccl" if use_cuda else "gloo", "ranks": out}), flush=True)
    dist.destroy_process_group()
    return 0 if ok else 1


def smoke(root: str, gpus: int = 0, port: int = 29431) -> dict:
    rep: dict = {"root": str(root)}
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=name,memory.total", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=30)
        rep["gpus"] = [l.strip() for l in out.stdout.splitlines() if l.strip()] if out.returncode == 0 else []
    except (OSError, subprocess.TimeoutExpired):
        rep["gpus"] = []
    rep["mounts"] = {}
    for m in MOUNTS:
        d = Path(root) / m
        d.mkdir(parents=True, exist_ok=True)
        probe = d / ".b2f-smoke"
        try:
            probe.write_text("ok")
            probe.unlink()
            rep["mounts"]["/" + m] = {"path": str(d), "writable": True}
        except OSError as e:
            rep["mounts"]["/" + m] = {"path": str(d), "writable": False, "error": str(e)}
    try:
        from ..ops import _lib

        _lib.host_lib()
        _lib.cuda_lib()
        rep["native_libs"] = _lib.loaded_libraries()
    except Exception as e:       # noqa: BLE001 - reported, not raised: this is a diagnosis command
        rep["native_libs"] = f"not loadable: {e}"
    world = gpus or max(len(rep["gpus"]), 2 if not rep["gpus"] else 1)
    rep["requested_ranks"] = world
    if gpus and rep["gpus"] and len(rep["gpus"]) < gpus:
        rep["collective"] = f"only {len(rep['gpus'])} of the {gpus} requested GPUs are visible"
        rep["ok"] = False
        return rep
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE=str(world))
    procs = [subprocess.Popen([sys.executable, "-m", "b200flow.launcher.tools"], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(world)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    rep["collective"] = next((json.loads(l[len("B2F_SMOKE "):]) for o in outs for l in o.splitlines()
                              if l.startswith("B2F_SMOKE ")), {"error": outs[0][-500:]})
    rep["ok"] = all(p.returncode == 0 for p in procs) and all(v["writable"] for v in rep["mounts"].values()) \
        and isinstance(rep["native_libs"], (list, dict))
    return rep


def stage_data(root: str, src: str, dst: str) -> dict:
    """Copy `src` (directory, file or s3://bucket/prefix) to `dst` (a pod path such as /efs/data/coco2017)."""
    if src.startswith("s3://"):
        s3_root = os.environ.get("B2F_S3_ROOT")
        if not s3_root:
            raise SystemExit("stage-data: no object store on this box; set B2F_S3_ROOT to a directory that "
                             "holds <bucket>/<prefix> trees, or pass a local source")
        src = str(Path(s3_root) / src[len("s3://"):])
    s = Path(src)
    if not s.exists():
        raise SystemExit(f"stage-data: source {s} does not exist")
    parts = Path(dst).parts
    if len(parts) < 2 or parts[0] != "/" or parts[1] not in MOUNTS:
        raise SystemExit("stage-data: destination must be under /efs or /fsx")
    d = Path(root).joinpath(*parts[1:])
    d.parent.mkdir(parents=True, exist_ok=True)
    if s.is_dir():
        shutil.copytree(s, d, dirs_exist_ok=True)
    else:
        d.mkdir(parents=True, exist_ok=True)
        shutil.copy2(s, d / s.name)
    files = [p for p in d.rglob("*") if p.is_file()]
    return {"staged_to": str(d), "files": len(files), "bytes": sum(p.stat().st_size for p in files)}


def attach(root: str, release: str | None = None, depth: int = 3) -> dict:
    out: dict = {}
    for m in MOUNTS:
        base = Path(root) / m
        tops = [base / "home" / release] if release else [base]
        if release and m == "fsx":
            tops.append(base / "logs")
        listing = []
        for top in tops:
            if top.exists():
                for p in sorted(top.rglob("*")):
                    rel = p.relative_to(base)
                    if len(rel.parts) <= depth + (2 if release else 0):
                        listing.append(("/" + m + "/" + str(rel)) + ("/" if p.is_dir() else ""))
        out["/" + m] = {"path": str(base), "entries": listing}
    return out


if __name__ == "__main__":
    sys.exit(_rank_main())
