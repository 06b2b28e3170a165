"""Hyper-parameter sweeps over the launcher, shaped like a Katib experiment.

The reference's platform installs Kubeflow Katib 0.15 (charts/machine-learning/kubeflow-platform: algorithms
`random` / `grid` / ..., trial templates with `${trialParameters.<name>}` placeholders, StdOut / File / TFEvent
metric collectors, median-stop early stopping — SURVEY.md §2A "P/kubeflow-katib").  SURVEY lists it as platform,
not as part of the build; this module keeps the capability for the one-box launcher: an experiment is a YAML
with Katib's field names, a trial is one chart release, the objective is read from the release's logs.

    objective:        {type: minimize | maximize, objectiveMetricName: loss, goal: 0.05}
    algorithm:        {algorithmName: random | grid, algorithmSettings: [{name: random_state, value: "7"}]}
    maxTrialCount: 8 ; parallelTrialCount: 2 ; maxFailedTrialCount: 2
    parameters:       [{name: lr, parameterType: double | int | categorical | discrete,
                        feasibleSpace: {min: "1e-4", max: "1e-2", step: "...", list: [...]}}]
    metricsCollectorSpec: {collector: {kind: StdOut | File}, source: {fileSystemPath: {path: ...},
                                                                       filter: {metricsFormat: [regex]}}}
    earlyStopping:    {algorithmName: medianstop, algorithmSettings: [{name: min_trials_required, value: "3"}]}
    trialTemplate:    {chart: <chart>, releasePrefix: trial, values: {...}, set: ["k=${trialParameters.lr}"]}

Placeholders are substituted in every string of `values` and `set`.  Metrics are `name=value` (Katib's default
StdOut format) taken from the pods' logs, or from a JSON-lines file when the collector kind is File; the last
value of the objective metric is the trial's result, the sequence of values feeds median stopping.
"""
from __future__ import annotations

import copy
import itertools
import json
import random
import re
import threading
from pathlib import Path
from typing import Any

from ..launcher import runner as R
from .runner import ChartHandler

_DEFAULT_FORMAT = r"([\w|-]+)\s*=\s*([+-]?\d*(?:\.\d+)?(?:[Ee][+-]?\d+)?)"


def _settings(block: dict | None) -> dict:
    return {s["name"]: s["value"] for s in (block or {}).get("algorithmSettings", []) or []}


def _space(p: dict) -> list | tuple:
    """Categorical / discrete / stepped spaces as a list; continuous ones as (lo, hi, is_int)."""
    fs, kind = p.get("feasibleSpace", {}), p.get("parameterType", "double")
    if kind in ("categorical", "discrete"):
        vals = list(fs["list"])
        return [float(v) for v in vals] if kind == "discrete" else vals
    lo, hi = float(fs["min"]), float(fs["max"])
    if fs.get("step"):
        step, out, v = float(fs["step"]), [], lo
        while v <= hi + 1e-12:
            out.append(int(round(v)) if kind == "int" else round(v, 12))
            v += step
        return out
    return (lo, hi, kind == "int")


def suggestions(spec: dict) -> list[dict]:
    """The trial assignments the experiment's algorithm proposes (at most maxTrialCount)."""
    params = spec.get("parameters", [])
    algo = (spec.get("algorithm") or {}).get("algorithmName", "random")
    n = int(spec.get("maxTrialCount", 1))
    spaces = {p["name"]: _space(p) for p in params}
    if algo == "grid":
        axes = []
        for name, sp in spaces.items():
            if isinstance(sp, tuple):
                lo, hi, is_int = sp
                sp = list(range(int(lo), int(hi) + 1)) if is_int else [lo, (lo + hi) / 2, hi]
            axes.append([(name, v) for v in sp])
        return [dict(combo) for combo in itertools.islice(itertools.product(*axes), n)]
    if algo != "random":
        raise ValueError(f"algorithm {algo!r}: only random and grid are built in")
    rng = random.Random(int(_settings(spec.get("algorithm")).get("random_state", 0)))
    out = []
    for _ in range(n):
        a = {}
        for name, sp in spaces.items():
            if isinstance(sp, tuple):
                lo, hi, is_int = sp
                a[name] = rng.randint(int(lo), int(hi)) if is_int else rng.uniform(lo, hi)
            else:
                a[name] = rng.choice(sp)
        out.append(a)
    return out


def _substitute(obj: Any, assignment: dict) -> Any:
    if isinstance(obj, str):
        return re.sub(r"\$\{trialParameters\.([\w-]+)\}", lambda m: str(assignment[m.group(1)]), obj)
    if isinstance(obj, dict):
        return {k: _substitute(v, assignment) for k, v in obj.items()}
    if isinstance(obj, list):
        return [_substitute(v, assignment) for v in obj]
    return obj


def collect_metrics(spec: dict, root: str, release: str) -> dict[str, list[float]]:
    """Every value of every metric the collector finds for a release, in log order."""
    mc = spec.get("metricsCollectorSpec") or {}
    kind = (mc.get("collector") or {}).get("kind", "StdOut")
    src = mc.get("source") or {}
    formats = (src.get("filter") or {}).get("metricsFormat") or [_DEFAULT_FORMAT]
    out: dict[str, list[float]] = {}
    if kind == "File":
        path = str((src.get("fileSystemPath") or {}).get("path", "")).replace("${release}", release)
        texts = [Path(p).read_text(errors="replace") for p in ([path] if Path(path).exists() else [])]
    else:
        texts = list(R.logs(root, release, None).values())
    for text in texts:
        for line in text.splitlines():
            line = line.strip()
            if line.startswith("{"):
                try:
                    for k, v in json.loads(line).items():
                        if isinstance(v, (int, float)) and not isinstance(v, bool):
                            out.setdefault(k, []).append(float(v))
                    continue
                except ValueError:
                    pass
            for fmt in formats:
                for name, val in re.findall(fmt, line):
                    try:
                        out.setdefault(name, []).append(float(val))
                    except ValueError:
                        pass
    return out


def run_sweep(spec: dict, root: str = "/tmp/b200flow", log=print) -> dict:
    obj = spec["objective"]
    metric, minimize = obj["objectiveMetricName"], obj.get("type", "minimize") == "minimize"
    goal = obj.get("goal")
    tmpl = spec["trialTemplate"]
    prefix = tmpl.get("releasePrefix", "trial")
    parallel = max(1, int(spec.get("parallelTrialCount", 1)))
    max_failed = int(spec.get("maxFailedTrialCount", 0))
    es = spec.get("earlyStopping") or {}
    min_trials = int(_settings(es).get("min_trials_required", 3)) if es.get("algorithmName") == "medianstop" else None
    trials, lock, state = [], threading.Lock(), {"failed": 0, "stop": False}
    queue = list(enumerate(suggestions(spec)))

    def better(a, b):
        return a < b if minimize else a > b

    def run_one(i, assignment):
        release = f"{prefix}-{i}"
        cfg = {"release_name": release, "chart": tmpl["chart"], "values": _substitute(copy.deepcopy(tmpl.get("values", {})), assignment),
               "set": _substitute(list(tmpl.get("set", [])), assignment), "max_restarts": tmpl.get("max_restarts", 0),
               "pod_complete_timeout": tmpl.get("timeout", 7 * 24 * 3600), "strict_exit": tmpl.get("strict_exit", True)}
        handler = ChartHandler(cfg, root, log=lambda m: None)
        rc = 1
        try:
            handler.install()
            rc = handler.wait()
        except Exception as e:  # noqa: BLE001
            log(f"{release}: {e!r}")
        values = collect_metrics(spec, root, release).get(metric, [])
        handler.uninstall()
        rec = {"name": release, "assignment": assignment, "exit_code": rc, "values": values,
               "objective": values[-1] if values else None, "status": "Succeeded" if rc == 0 and values else "Failed"}
        with lock:
            if rec["status"] == "Failed":
                state["failed"] += 1
                if state["failed"] > max_failed:
                    state["stop"] = True
            done = [t for t in trials if t["status"] == "Succeeded"]
            if rec["status"] == "Succeeded" and min_trials is not None and len(done) >= min_trials:
                # median stopping rule, applied after the fact on one box: a trial whose best value is worse than
                # the median of the other trials' running averages is marked as one Katib would have stopped early
                med = sorted(sum(t["values"]) / len(t["values"]) for t in done)[len(done) // 2]
                best_seen = min(values) if minimize else max(values)
                if better(med, best_seen):
                    rec["status"] = "EarlyStopped"
            trials.append(rec)
            if goal is not None and rec["objective"] is not None and (
                    rec["objective"] <= float(goal) if minimize else rec["objective"] >= float(goal)):
                state["stop"] = True                       # goal reached: no further trials are started
        log(f"{release}: {assignment} -> {metric}={rec['objective']} ({rec['status']})")

    def worker():
        while True:
            with lock:
                if state["stop"] or not queue:
                    return
                i, a = queue.pop(0)
            run_one(i, a)

    threads = [threading.Thread(target=worker) for _ in range(parallel)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    ok = [t for t in trials if t["objective"] is not None and t["status"] != "Failed"]
    best = (min if minimize else max)(ok, key=lambda t: t["objective"]) if ok else None
    status = "Failed" if state["failed"] > max_failed else ("Succeeded" if best else "Failed")
    return {"status": status, "objective": metric, "best": best, "trials": sorted(trials, key=lambda t: t["name"])}
