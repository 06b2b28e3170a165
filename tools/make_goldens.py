"""Regenerate tests/golden/: a fingerprint (SHA-256, line count) of the rendered job script plus the
launch plan and environment of every reference values file (fixed release name, time and root), which
tests/test_jobspec_golden.py compares against.  Fingerprints rather than the scripts themselves: the
scripts are mostly the reference's own pre_script lines, which do not belong in this repository."""
import datetime as dt
import hashlib
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from b200flow.jobspec.render import render_job  # noqa: E402

EX = Path("/root/reference/examples")
T0 = dt.datetime(2026, 1, 2, 3, 4, 5)
CASES = [
    ("pytorchjob-elastic", "accelerate/bert-glue-mrpc/pretrain.yaml"),
    ("pytorchjob-distributed", "megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml"),
    ("pytorchjob-distributed", "megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml"),
    ("data-process", "megatron-deepspeed/gpt2_345m/wikicorpus.yaml"),
    ("raytrain", "ray/lightning-bert/fine-tune.yaml"),
    ("mpijob-horovod-tensorflow-gpu", "maskrcnn/train-maskrcnn-tensorpack.yaml"),
    ("mpijob-horovod-tensorflow-gpu", "maskrcnn/train-maskrcnn-aws.yaml"),
]


def golden_name(chart, rel):
    return f"{chart}__{rel.replace('/', '__').removesuffix('.yaml')}"


def render_case(chart, rel):
    plan = render_job(chart, "golden", [EX / rel], root=Path("/b2f-root"), when=T0)
    meta = {"script_sha256": hashlib.sha256(plan.script.encode()).hexdigest(),
            "script_lines": len(plan.script.splitlines()),
            "kind": plan.kind, "nnodes": plan.nnodes, "nproc_per_node": plan.nproc_per_node,
            "env": dict(sorted(plan.env.items())),
            "mpirun_args": list(getattr(plan, "mpirun_args", []) or [])}
    return plan.script, meta


if __name__ == "__main__":
    out = ROOT / "tests" / "golden"
    out.mkdir(exist_ok=True)
    for chart, rel in CASES:
        script, meta = render_case(chart, rel)
        (out / (golden_name(chart, rel) + ".json")).write_text(json.dumps(meta, indent=1, sort_keys=True) + "\n")
        print("wrote", golden_name(chart, rel))
