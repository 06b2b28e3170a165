"""`python -m b200flow.pipeline run chart_configs.yaml [--root DIR]`      sequential chart pipeline (KFP component)
`python -m b200flow.pipeline sweep experiment.yaml [--root DIR]`        Katib-shaped hyper-parameter sweep"""
import argparse
import json
import sys

from ..jobspec.values import load_yaml
from .runner import run_pipeline

ap = argparse.ArgumentParser(prog="b200flow-pipeline")
ap.add_argument("cmd", choices=["run", "sweep"])
ap.add_argument("configs")
ap.add_argument("--root", default="/tmp/b200flow")
a = ap.parse_args()
data = load_yaml(a.configs)
if a.cmd == "sweep":
    from .sweep import run_sweep

    spec = data.get("spec", data) if isinstance(data, dict) else data      # a Katib Experiment object or its spec
    res = run_sweep(spec, a.root)
    print(json.dumps(res, indent=1, default=str))
    sys.exit(0 if res["status"] == "Succeeded" else 1)
res = run_pipeline(data["chart_configs"] if isinstance(data, dict) else data, a.root)
print(res)
sys.exit(0 if res == "Success" else 1)
