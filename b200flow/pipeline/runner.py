"""Sequential chart pipeline with the semantics of the reference's KFP component.

`helm_charts_component(chart_configs) -> "Success" | "Failure"`
(kfp/components/src/helm-charts-component/helm_charts_component.py:199-205): for every
config, install the chart, wait for its pods, and ALWAYS uninstall afterwards (`:29-40`);
stop at the first failure (`:202-203`); SIGINT / SIGTERM / SIGABRT uninstall the running
release before exiting (`:22-27`).  Pod discovery by annotations and phase polling collapse
here into the runner's own supervision, and `wait_for_pods` (`:135-197`) keeps its shape: the
release's phase is read every `pod_check_secs` (default 300 s, `:172`) while it is Running and
more often while it is Pending; `pod_error_timeout` bounds the time it may stay Pending (waiting
for GPUs that other releases hold; default 1800 s), `pod_complete_timeout` the time it may run
(default 7 days, `:142-143`); Failed / Unknown fail the pipeline, a vanished release counts as
Succeeded (`:181-183`).

Config keys: release_name, chart (name or path), values (dict) / values_files, set, timeout,
pod_complete_timeout, pod_error_timeout, pod_check_secs, strict_exit.  `repo_url`/`path`/`branch`/`version` are
accepted and only used to derive the chart name.
"""
from __future__ import annotations

import signal
import threading
import time
from typing import Any

from ..jobspec.render import render_job
from ..launcher import runner as R


class ChartHandler:
    def __init__(self, cfg: dict, root: str, log=print):
        self.cfg, self.root, self.log = cfg, root, log
        self.release = cfg["release_name"]
        self.runner: R.Runner | None = None
        self._old = {}

    def _install_signals(self):
        for s in (signal.SIGINT, signal.SIGTERM, signal.SIGABRT):
            try:
                self._old[s] = signal.signal(s, self._on_signal)
            except ValueError:
                pass

    def _restore_signals(self):
        for s, h in self._old.items():
            signal.signal(s, h)

    def _on_signal(self, signum, frame):
        self.log(f"signal {signum}: uninstalling {self.release}")
        self.uninstall()
        raise SystemExit(128 + signum)

    def install(self) -> None:
        cfg = self.cfg
        chart = cfg.get("chart") or cfg.get("path") or ""
        vals = [cfg["values"]] if isinstance(cfg.get("values"), dict) else []
        vals += list(cfg.get("values_files") or [])
        plan = render_job(chart, self.release, vals, cfg.get("set") or [], root=self.root,
                          strict_exit=bool(cfg.get("strict_exit", False)))
        self.runner = R.Runner(plan, self.root, offline=not cfg.get("online", False),
                               max_restarts=cfg.get("max_restarts"))
        self.log(f"installed {self.release} ({plan.chart}, {plan.kind})")

    def wait(self) -> int:
        complete_timeout = float(self.cfg.get("pod_complete_timeout", 7 * 24 * 3600))
        error_timeout = float(self.cfg.get("pod_error_timeout", 1800))
        check = float(self.cfg.get("pod_check_secs", 300))
        out: dict = {}

        def supervise():
            try:
                out["rc"] = self.runner.run(timeout=complete_timeout, pending_timeout=error_timeout)
            except BaseException as e:  # noqa: BLE001 - reported as phase Unknown below
                out["error"] = e

        th = threading.Thread(target=supervise, name=f"release-{self.release}", daemon=True)
        th.start()
        start = time.time()
        phase = "Pending"
        while th.is_alive():
            th.join(timeout=check if phase == "Running" else min(check, 60.0, 1.0 if phase == "Pending" else 60.0))
            st = R.status(self.root, self.release)
            phase = st.get("phase") or "Succeeded"           # status gone: the reference treats Not Found as done
            if phase in ("Restarting",):
                phase = "Running"
            self.log(f"release {self.release} phase: {phase} ({time.time() - start:.0f} s)")
        if "error" in out:
            self.log(f"{self.release}: phase Unknown ({out['error']!r})")
            return 1
        return int(out.get("rc", 1))

    def uninstall(self) -> None:
        if self.runner is not None:
            self.runner.stop()
        R.uninstall(self.root, self.release)
        self.log(f"uninstalled {self.release}")

    def __call__(self) -> int:
        self._install_signals()
        try:
            self.install()
            return self.wait()
        except Exception as e:  # noqa: BLE001
            self.log(f"{self.release}: {e!r}")
            return 1
        finally:
            self.uninstall()
            self._restore_signals()


def run_pipeline(chart_configs: list[dict[str, Any]], root: str = "/tmp/b200flow", log=print) -> str:
    for cfg in chart_configs:
        if ChartHandler(cfg, root, log)() != 0:
            return "Failure"
    return "Success"
