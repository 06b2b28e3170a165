"""Ray-Train-shaped worker group: failure budget with restart from the newest reported checkpoint, explicit resume,
metrics history (b200flow/engines/ray_train_api.py)."""
import json
import os
import sys
import tempfile
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.engines import ray_train_api as train  # noqa: E402
from b200flow.engines.ray_train_api import (Checkpoint, CheckpointConfig, FailureConfig, RunConfig,  # noqa: E402
                                            ScalingConfig, TorchTrainer)


def _loop(config):
    """Counts epochs; continues from a checkpoint when there is one; crashes once at epoch 2 if asked to."""
    ctx = train.get_context()
    start = 0
    ck = train.get_checkpoint()
    if ck is not None:
        with ck.as_directory() as d:
            start = json.loads((Path(d) / "state.json").read_text())["epoch"] + 1
    for epoch in range(start, config["epochs"]):
        marker = Path(config["marker"])
        if config.get("crash_at") == epoch and ctx.get_world_rank() == 1 and not marker.exists():
            marker.write_text("crashed once")
            raise RuntimeError("injected worker failure")
        tmp = tempfile.mkdtemp()
        (Path(tmp) / "state.json").write_text(json.dumps({"epoch": epoch}))
        train.report({"epoch": epoch, "started_from": start}, checkpoint=Checkpoint.from_directory(tmp))


def _trainer(tmp_path, max_failures, crash_at=None, resume=None, epochs=4):
    return TorchTrainer(_loop, train_loop_config={"epochs": epochs, "crash_at": crash_at, "marker": str(tmp_path / "marker")},
                        scaling_config=ScalingConfig(num_workers=2, use_gpu=False),
                        run_config=RunConfig(name="run", storage_path=str(tmp_path), failure_config=FailureConfig(max_failures),
                                             checkpoint_config=CheckpointConfig(num_to_keep=None)),
                        resume_from_checkpoint=resume)


def test_worker_failure_restarts_the_group_from_the_newest_checkpoint(tmp_path):
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    res = _trainer(tmp_path, max_failures=1, crash_at=2).fit()
    epochs = [m["epoch"] for m in res.metrics_history]
    assert epochs == [0, 1, 2, 3]                                       # nothing repeated, nothing lost
    # rank 0 had already reported epoch 2 (with its checkpoint) when rank 1 failed in that epoch: the restart begins at 3
    assert [m["started_from"] for m in res.metrics_history] == [0, 0, 0, 3]
    assert res.metrics == {"epoch": 3, "started_from": 3} and res.error is None
    assert list(res.metrics_dataframe["epoch"]) == [0, 1, 2, 3]
    assert len((tmp_path / "run" / "result.json").read_text().splitlines()) == 4
    assert sorted(p.name for p in (tmp_path / "run").glob("checkpoint_*")) == [f"checkpoint_{i:06d}" for i in range(4)]


def test_failure_without_budget_raises_and_explicit_resume_continues(tmp_path):
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    with pytest.raises(RuntimeError, match="injected worker failure"):
        _trainer(tmp_path, max_failures=0, crash_at=1).fit()
    ck = Checkpoint(str(tmp_path / "run" / "checkpoint_000000"))
    res = _trainer(tmp_path, max_failures=0, resume=ck).fit()           # the marker exists: no second crash
    assert [m["epoch"] for m in res.metrics_history] == [1, 2, 3] and res.metrics["started_from"] == 1
