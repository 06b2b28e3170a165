"""The parts of the bench.py / __graft_entry__.py contract that can be checked without a GPU."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_reports_unavailable_and_exits_zero():
    res = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--gpus", "1", "--steps", "2", "--warmup", "1"],
                         cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert res.returncode == 0
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and isinstance(line["unavailable"], str) and "\n" not in line["unavailable"]


def test_defaults_and_entry_points():
    sys.path.insert(0, str(ROOT))
    import importlib

    bench = importlib.import_module("bench")
    old = sys.argv
    sys.argv = ["bench.py"]
    try:
        a = bench.parse_args()
    finally:
        sys.argv = old
    assert a.gpus == 1 and a.warmup >= 3 and a.steps >= 1 and a.impl == "b200flow" and a.workload == "llama2-7b"
    assert bench.topology_for(8) == (2, 2) and bench.topology_for(1) == (1, 1)        # TP2 x PP2 (x DP2) on 8 GPUs
    entry = importlib.import_module("__graft_entry__")
    assert callable(entry.build) and callable(entry.smoke)
    src = (ROOT / "bench.py").read_text()
    for key in ('"metric"', '"value"', '"n_gpus"', '"ms_per_step"', '"higher_is_better"', '"scaling"', '"vs_baseline"',
                '"dtype"', '"data"', '"config"', '"clocks"', '"e2e"', '"h2d_bytes_per_step"', '"d2h_bytes_per_step"',
                '"gpu_launches"'):
        assert key in src, key
