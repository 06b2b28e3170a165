"""Sequential chart pipeline with the semantics of the reference's KFP component.

`helm_charts_component(chart_configs) -> "Success" | "Failure"`
(kfp/components/src/helm-charts-component/helm_charts_component.py:199-205): for every
config, install the chart, wait for its pods, and ALWAYS uninstall afterwards (`:29-40`);
stop at the first failure (`:202-203`); SIGINT / SIGTERM / SIGABRT uninstall the running
release before exiting (`:22-27`).  Pod discovery by annotations and phase polling collapse
here into the runner's own supervision; the timeouts keep their names:
`pod_error_timeout` bounds the time a release may stay Pending, `pod_complete_timeout` the
time it may run (defaults 1800 s and 7 days, `:142-143`).

Config keys: release_name, chart (name or path), values (dict) / values_files, set, timeout,
pod_complete_timeout, pod_error_timeout, strict_exit.  `repo_url`/`path`/`branch`/`version` are
accepted and only used to derive the chart name.
"""
from __future__ import annotations

import signal
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
        timeout = float(self.cfg.get("pod_complete_timeout", 7 * 24 * 3600))
        return self.runner.run(timeout=timeout)

    def uninstall(self) -> None:
        if self.runner is not None:
            self.runner._kill_all()
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
