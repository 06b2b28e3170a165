"""Golden tests of the job-spec layer (plan item 1): every reference values file is
rendered with a fixed release name / time / root and compared with the committed fingerprint of the
script and the full launch plan (kind, world shape, environment, mpirun arguments).  Regenerate with
`python tools/make_goldens.py` after an intended change."""
import json
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
import make_goldens as G  # noqa: E402

pytestmark = pytest.mark.skipif(not G.EX.exists(), reason="reference tree not mounted")


@pytest.mark.parametrize("chart,rel", G.CASES)
def test_rendering_matches_golden(chart, rel):
    script, meta = G.render_case(chart, rel)
    want = json.loads((ROOT / "tests/golden" / (G.golden_name(chart, rel) + ".json")).read_text())
    assert meta["kind"] == want["kind"]
    assert (meta["nnodes"], meta["nproc_per_node"]) == (want["nnodes"], want["nproc_per_node"])
    assert meta["env"] == want["env"]
    assert meta["mpirun_args"] == want["mpirun_args"]
    assert meta["script_lines"] == want["script_lines"]
    assert meta["script_sha256"] == want["script_sha256"], "rendered script changed: diff it, then tools/make_goldens.py"
    # rendering is a pure function of (chart, values, release, time, root)
    assert G.render_case(chart, rel)[0] == script
