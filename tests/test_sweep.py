"""Katib-shaped hyper-parameter sweeps over the launcher (b200flow/pipeline/sweep.py)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.pipeline.sweep import run_sweep, suggestions  # noqa: E402

SPEC = {
    "objective": {"type": "minimize", "objectiveMetricName": "loss"},
    "algorithm": {"algorithmName": "grid"},
    "maxTrialCount": 6, "parallelTrialCount": 2, "maxFailedTrialCount": 1,
    "parameters": [
        {"name": "lr", "parameterType": "double", "feasibleSpace": {"min": "0.1", "max": "0.5", "step": "0.2"}},
        {"name": "opt", "parameterType": "categorical", "feasibleSpace": {"list": ["adam", "sgd"]}},
    ],
    "trialTemplate": {
        "chart": "data-process", "releasePrefix": "hpo",
        "values": {"process": {"command": ["python3", "-c"],
                               "args": ["'print(\"step=1 loss=9.0\"); "
                                        "print(\"loss=%.4f\" % ((${trialParameters.lr} - 0.3) ** 2 + "
                                        "(0.5 if \"${trialParameters.opt}\" == \"sgd\" else 0.0)))'"]}},
    },
}


def test_grid_and_random_suggestions():
    grid = suggestions(SPEC)
    assert len(grid) == 6 and {(round(g["lr"], 1), g["opt"]) for g in grid} == {
        (lr, o) for lr in (0.1, 0.3, 0.5) for o in ("adam", "sgd")}
    rnd = dict(SPEC, algorithm={"algorithmName": "random", "algorithmSettings": [{"name": "random_state", "value": "3"}]},
               maxTrialCount=20, parameters=[
                   {"name": "lr", "parameterType": "double", "feasibleSpace": {"min": "0.001", "max": "0.01"}},
                   {"name": "layers", "parameterType": "int", "feasibleSpace": {"min": "2", "max": "5"}}])
    a, b = suggestions(rnd), suggestions(rnd)
    assert a == b and len(a) == 20                                         # seeded: reproducible
    assert all(0.001 <= t["lr"] <= 0.01 and t["layers"] in (2, 3, 4, 5) for t in a)
    assert len({t["layers"] for t in a}) > 1


def test_sweep_runs_trials_as_chart_releases_and_picks_the_best(tmp_path):
    res = run_sweep(SPEC, root=str(tmp_path), log=lambda m: None)
    assert res["status"] == "Succeeded" and len(res["trials"]) == 6
    best = res["best"]
    assert round(best["assignment"]["lr"], 1) == 0.3 and best["assignment"]["opt"] == "adam" and best["objective"] == 0.0
    t = next(t for t in res["trials"] if t["assignment"]["opt"] == "sgd" and round(t["assignment"]["lr"], 1) == 0.1)
    assert t["values"] == [9.0, 0.54] and t["status"] == "Succeeded"      # every reported value is kept, the last one counts
    # a goal stops the experiment early; failing trials beyond the budget fail it
    early = run_sweep(dict(SPEC, objective={"type": "minimize", "objectiveMetricName": "loss", "goal": 0.05},
                           parallelTrialCount=1), root=str(tmp_path / "goal"), log=lambda m: None)
    assert early["status"] == "Succeeded" and len(early["trials"]) < 6 and early["best"]["objective"] <= 0.05
    bad = dict(SPEC, maxTrialCount=3, maxFailedTrialCount=0, parallelTrialCount=1)
    bad["trialTemplate"] = {"chart": "data-process", "values": {"process": {"command": ["false"], "args": []}}}
    failed = run_sweep(bad, root=str(tmp_path / "bad"), log=lambda m: None)
    assert failed["status"] == "Failed" and len(failed["trials"]) == 1
