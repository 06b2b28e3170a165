"""Single-GPU run of the fused compute+collective kernels (K1-K5) in loopback: rank 0 of a world of
two with the peer's symmetric buffers on the same device and the peer's flags pre-set
(tools/prof_fused_loopback.py).  Exercises the bulk-copy gather role, staged bulk sends, flag
protocol and both GEMM cores without needing a second GPU."""
import os
import re
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    res = subprocess.run([sys.executable, str(ROOT / "tools" / "prof_fused_loopback.py")], capture_output=True,
                         text=True, timeout=300, cwd=ROOT, env=env)
    out = res.stdout + res.stderr
    assert res.returncode == 0, out[-1500:]
    errs = {m.group(1): float(m.group(2)) for m in re.finditer(r"(K3|K4) loopback rel err ([0-9.e+-]+)", out)}
    assert set(errs) == {"K3", "K4"}, out[-800:]
    assert all(v < 2e-2 for v in errs.values()), errs
    ar = [float(m.group(1)) for m in re.finditer(r"K4AR loopback rel err ([0-9.e+-]+)", out)]
    assert len(ar) == 2 and all(v < 2e-2 for v in ar), out[-800:]
    assert "K1/K2 loopback done" in out
    m = re.search(r"K5 loopback max \|combine\(dispatch\(x\)\) - x\| ([0-9.e+-]+)", out)
    assert m and float(m.group(1)) == 0.0, out[-400:]


def test_fused_kernels_loopback_cta_pair_core():
    _run({})


def test_fused_kernels_loopback_single_cta_core():
    _run({"B2F_FUSED_TP_CORE": "1"})
