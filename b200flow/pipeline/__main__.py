"""`python -m b200flow.pipeline run chart_configs.yaml [--root DIR]`"""
import argparse
import sys

from ..jobspec.values import load_yaml
from .runner import run_pipeline

ap = argparse.ArgumentParser(prog="b200flow-pipeline")
ap.add_argument("cmd", choices=["run"])
ap.add_argument("configs")
ap.add_argument("--root", default="/tmp/b200flow")
a = ap.parse_args()
data = load_yaml(a.configs)
res = run_pipeline(data["chart_configs"] if isinstance(data, dict) else data, a.root)
print(res)
sys.exit(0 if res == "Success" else 1)
