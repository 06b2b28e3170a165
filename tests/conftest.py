"""pytest configuration: registers the `gpu` marker and shared fixtures.

`python -m pytest tests -m "not gpu"` runs everywhere (CPU/gloo);
`python -m pytest tests -m gpu` needs a B200 and the in-tree CUDA library.
"""
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200, sm_100a)")
    config.addinivalue_line("markers", "multigpu: test needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:  # pragma: no cover
        ngpu = 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and ngpu == 0:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture
def tmp_root(tmp_path):
    """A scratch root holding local stand-ins for the /efs and /fsx mounts."""
    (tmp_path / "efs").mkdir()
    (tmp_path / "fsx").mkdir()
    return tmp_path
