"""`python -m b200flow.launcher` — helm-like CLI over the local job runner.

    install   <release> <chart> [-f values.yaml ...] [--set k=v ...] [--root DIR] [--wait/--no-wait]
    template  <release> <chart> [-f ...] [--set ...]        (print the rendered script + plan)
    uninstall <release> [--root DIR]
    status    <release> [--root DIR]
    logs      <release> [--root DIR] [--tail N] [-f]
    list      [--root DIR]                       (releases with chart, phase, restarts)
    scale     <release> <replicas> | --replicas N [--root DIR]  (elastic releases: worker pods within min_replicas..max_replicas)
    smoke     [--gpus N] [--root DIR]            (GPU / mounts / native libraries / collective check)
    stage-data <src> <dst under /efs or /fsx> [--root DIR]
    attach    [release] [--root DIR]             (where /efs and /fsx live locally and what is in them)

Mirrors `helm install <release> <chart> -f <values.yaml> -n <ns>`
(examples/accelerate/bert-glue-mrpc/README.md:23-25).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from dataclasses import asdict

from ..jobspec.render import render_job
from . import runner as R


def _common(p):
    p.add_argument("--root", default=os.environ.get("B2F_ROOT", "/tmp/b200flow"))
    # accepted so that the README command lines of the reference work with `helm` / `kubectl` swapped for this CLI
    # (`helm install --debug <release> ./charts/machine-learning/training/<chart>/ -f <values> -n <namespace>`,
    # examples/accelerate/bert-glue-mrpc/README.md:23-25); there is one namespace here
    p.add_argument("--debug", action="store_true", help="accepted for `helm install --debug`; prints the launch plan")
    if not any(a.dest == "namespace" for a in p._actions):
        p.add_argument("-n", "--namespace", default="local", help="accepted for helm / kubectl command lines; unused")


def _install_detached(a, plan, argv) -> int:
    """`--no-wait`: a supervisor process of its own runs the release; this call returns once the pods are up."""
    import subprocess
    import time
    from pathlib import Path

    state = R.ReleaseState(Path(a.root), a.release)
    state.dir.mkdir(parents=True, exist_ok=True)
    (state.dir / "status.json").unlink(missing_ok=True)
    cmd = [sys.executable, "-m", "b200flow.launcher"] + [x for x in argv if x != "--no-wait"]
    log = open(state.dir / "supervisor.log", "ab")
    sup = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, start_new_session=True)
    deadline = time.time() + 30
    while time.time() < deadline and sup.poll() is None and state.read_status().get("phase") is None:
        time.sleep(0.1)
    if state.exists():
        state.write_status(supervisor_pid=sup.pid)
    print(json.dumps(R.status(a.root, a.release)))
    return 0 if sup.poll() in (None, 0) else int(sup.returncode)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="b200flow")
    sub = ap.add_subparsers(dest="cmd", required=True)
    for name in ("install", "template"):
        p = sub.add_parser(name)
        p.add_argument("release")
        p.add_argument("chart")
        p.add_argument("-f", "--values", action="append", default=[])
        p.add_argument("--set", action="append", default=[], dest="sets")
        p.add_argument("-n", "--namespace", default="local")
        p.add_argument("--strict-exit", action="store_true",
                       help="propagate the training command's exit status (reference masks it with | tee)")
        p.add_argument("--online", action="store_true", help="do not install the offline git/pip shims")
        p.add_argument("--timeout", type=float, default=None)
        p.add_argument("--no-wait", action="store_true",
                       help="return as soon as the pods run, like `helm install` without --wait; follow with status / logs")
        p.add_argument("--wait", action="store_true", help="wait for the job to finish (the default)")
        p.add_argument("--inject-fault", default=None, help="rank=R,step=S,kind={exit,hang,sigkill}")
        p.add_argument("--max-restarts", type=int, default=None)
        p.add_argument("--hang-timeout", type=float, default=None,
                       help="seconds of total log silence after which the process group is killed and restarted")
        p.add_argument("--auto-resume", action="store_true",
                       help="restarted (or re-installed) workloads continue from their newest checkpoint even though the "
                            "values file passes no --load / --resume_from_checkpoint (the reference's examples do not)")
        _common(p)
    for name in ("uninstall", "status", "logs"):
        p = sub.add_parser(name)
        p.add_argument("release")
        p.add_argument("--tail", type=int, default=None)
        p.add_argument("-f", "--follow", action="store_true", help="logs: keep printing until the release stops running")
        _common(p)
    p = sub.add_parser("list")
    _common(p)
    p = sub.add_parser("scale")
    p.add_argument("release")
    p.add_argument("replicas_pos", type=int, nargs="?", default=None, metavar="replicas")
    p.add_argument("--replicas", type=int, default=None, help="like `kubectl scale --replicas N`")
    _common(p)
    p = sub.add_parser("smoke")
    p.add_argument("--gpus", type=int, default=0)
    _common(p)
    p = sub.add_parser("stage-data")
    p.add_argument("src")
    p.add_argument("dst")
    _common(p)
    p = sub.add_parser("attach")
    p.add_argument("release", nargs="?", default=None)
    _common(p)
    a = ap.parse_args(argv)
    if a.cmd == "list":        # `helm list`: every release under the root with its chart, phase and restart count
        from pathlib import Path

        rows = []
        for d in sorted((Path(a.root) / ".b200flow" / "releases").glob("*")):
            st = R.status(a.root, d.name)
            plan = json.loads((d / "plan.json").read_text()) if (d / "plan.json").exists() else {}
            rows.append({"name": d.name, "chart": plan.get("chart"), "kind": plan.get("kind"), "phase": st.get("phase"),
                         "restarts": st.get("restarts", 0), "pods": len(st.get("pods") or {})})
        print(json.dumps(rows, indent=1))
        return 0
    if a.cmd == "scale":
        n = a.replicas if a.replicas is not None else a.replicas_pos
        if n is None:
            ap.error("scale: give the number of replicas (positional or --replicas N)")
        ok = R.scale(a.root, a.release, n)
        print(f'release "{a.release}" ' + (f"scaling to {n}" if ok else "not found"))
        return 0 if ok else 1
    if a.cmd in ("smoke", "stage-data", "attach"):
        from . import tools as T

        if a.cmd == "smoke":
            rep = T.smoke(a.root, a.gpus)
            print(json.dumps(rep, indent=1))
            return 0 if rep.get("ok") else 1
        print(json.dumps(T.stage_data(a.root, a.src, a.dst) if a.cmd == "stage-data" else T.attach(a.root, a.release),
                         indent=1))
        return 0

    if a.cmd in ("install", "template"):
        plan = render_job(a.chart, a.release, a.values, a.sets, root=a.root, strict_exit=a.strict_exit)
        if a.cmd == "template":
            d = asdict(plan)
            script = d.pop("script")
            print(json.dumps(d, indent=1, default=str))
            print("---")
            print(script)
            return 0
        if a.debug:
            d = asdict(plan)
            d.pop("script")
            print(json.dumps(d, indent=1, default=str), file=sys.stderr)
        if a.no_wait:
            return _install_detached(a, plan, argv if argv is not None else sys.argv[1:])
        run = R.Runner(plan, a.root, offline=not a.online, fault=a.inject_fault,
                       max_restarts=a.max_restarts,
                       extra_env={"B2F_AUTO_RESUME": "1"} if a.auto_resume else None)
        rc = run.run(timeout=a.timeout, hang_timeout=a.hang_timeout)
        print(json.dumps(R.status(a.root, a.release)))
        return rc
    if a.cmd == "uninstall":
        ok = R.uninstall(a.root, a.release)
        print(f'release "{a.release}" ' + ("uninstalled" if ok else "not found"))
        return 0 if ok else 1
    if a.cmd == "status":
        print(json.dumps(R.status(a.root, a.release), indent=1))
        return 0
    if a.cmd == "logs":
        # `kubectl logs -f pytorchjob-<release>-worker-0` names a pod: "<kind>-<release>-<role>-<n>" of the operators
        # (examples/accelerate/bert-glue-mrpc/README.md:29); resolve it to the release and keep only that pod
        only = None
        if not R.logs(a.root, a.release):
            import re
            from pathlib import Path

            m = re.match(r"^(?:pytorchjob|mpijob|rayjob|data-process|coco-data|redpajama-data)-(.+?)(?:-(master|worker|launcher|head)-(\d+))?$",
                         a.release)
            if m and (Path(a.root) / ".b200flow" / "releases" / m.group(1)).is_dir():
                a.release = m.group(1)
                only = f"{m.group(1)}-{m.group(2)}-{m.group(3)}" if m.group(2) else None
        _logs = R.logs
        if only is not None:
            R_logs = lambda root, rel, tail=None: {k: v for k, v in _logs(root, rel, tail).items() if k.startswith(only)}  # noqa: E731
        else:
            R_logs = _logs
        for pod, txt in R_logs(a.root, a.release, a.tail).items():
            print(f"==> {pod} <==")
            print(txt)
        if a.follow:          # `kubectl logs -f`: new lines of every pod until the release is no longer running
            import time

            seen = {pod: len(txt) for pod, txt in R_logs(a.root, a.release).items()}
            while True:
                running = R.status(a.root, a.release).get("phase") in ("Running", "Restarting")
                for pod, txt in R_logs(a.root, a.release).items():
                    if len(txt) > seen.get(pod, 0):
                        for line in txt[seen.get(pod, 0):].splitlines():
                            print(f"[{pod}] {line}", flush=True)
                        seen[pod] = len(txt)
                if not running:
                    break
                time.sleep(0.5)
        return 0
    return 2


if __name__ == "__main__":
    sys.exit(main())
