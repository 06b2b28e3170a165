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


def test_stock_arm_is_attached_with_both_numbers_and_loss_parity():
    """`attach_baseline`: the stock arm's number and clocks travel in the line, `vs_baseline` is the ratio, and for the
    NCCL-collective arm (same engine, seed, data) the first three losses are held to 1e-2 relative."""
    sys.path.insert(0, str(ROOT))
    import importlib

    bench = importlib.import_module("bench")
    out = {"value": 100.0, "loss_curve": [11.19, 11.20, 3.79, 0.5]}
    base = {"impl": "b200flow-nccl-collectives", "value": 95.0, "unit": "tokens/s", "steps": 10, "ms_per_step": 1.0,
            "clocks": {"sm_mhz": 1500.0}, "e2e": {"value": 94.0}, "loss_curve": [11.19, 11.20, 3.80, 0.1]}
    bench.attach_baseline(out, base, 2)
    assert abs(out["vs_baseline"] - 100.0 / 95.0) < 1e-12
    assert out["baseline"]["value"] == 95.0 and out["baseline"]["e2e_value"] == 94.0 and "NCCL" in out["baseline"]["what"]
    assert out["loss_parity"]["ok"] and out["loss_parity"]["steps_compared"] == 3     # step 4 (0.5 vs 0.1) is not compared
    bad = {"value": 100.0, "loss_curve": [11.19, 12.5, 3.79]}
    bench.attach_baseline(bad, base, 2)
    assert not bad["loss_parity"]["ok"]
    twin = {"value": 50.0, "loss_curve": [11.0]}
    bench.attach_baseline(twin, {"impl": "twin", "value": 40.0, "loss_curve": [9.0]}, 1)
    assert twin["vs_baseline"] == 1.25 and "loss_parity" not in twin                 # a different model code: no parity claim
    failed = {"value": 1.0}
    bench.attach_baseline(failed, {"error": "stock arm exit 1"}, 2)
    assert failed["baseline"] == {"error": "stock arm exit 1"} and "vs_baseline" not in failed
