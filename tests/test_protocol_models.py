"""Exhaustive interleaving checks of the buffer-reuse protocols of K1, K2, K3, K4 and K5 (tools/protocol_model.py): with the
buffering the kernels use every read returns its own call's data under every schedule and nobody waits forever; the
mutated protocols (no closing barrier / one slot / one receive buffer / flags without the call counter) are caught."""
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))
import protocol_model as PM  # noqa: E402


@pytest.mark.parametrize("make", [PM.k1_model, PM.k2_model, PM.k3_model, PM.k4_model, PM.k5_model])
@pytest.mark.parametrize("world,calls", [(2, 4), (3, 3)])
def test_protocol_is_safe_under_every_interleaving(make, world, calls):
    assert PM.explore(*make(world, calls)) > 50


@pytest.mark.parametrize("make,mutation", [(PM.k1_model, {"end_barrier": False}), (PM.k2_model, {"start_barrier": False}), (PM.k3_model, {"slots": 1}), (PM.k4_model, {"buffers": 1}),
                                           (PM.k5_model, {"epoch_flags": False})])
def test_mutated_protocol_is_caught(make, mutation):
    with pytest.raises(PM.Violation):
        PM.explore(*make(2, 3, **mutation))


def test_deadlock_is_reported():
    # a program whose only step is never enabled
    with pytest.raises(PM.Violation, match="deadlock"):
        PM.explore(0, lambda s: iter(()), lambda s: False)
