"""The repository-named (hyphenated) package directory forwards to ``b200flow``."""
import importlib.util
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PK = ROOT / "amazon-eks-machine-learning-with-terraform-and-kubeflow_b200"


def test_alias_package_forwards():
    spec = importlib.util.spec_from_file_location("eks_ml_b200", PK / "__init__.py",
                                                  submodule_search_locations=[str(PK)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["eks_ml_b200"] = mod
    spec.loader.exec_module(mod)
    from eks_ml_b200.ops import gemm as g1
    from eks_ml_b200.models import gpt
    from eks_ml_b200.parallel import groups
    from eks_ml_b200.utils import flat
    import b200flow.ops.gemm as g2
    assert g1 is g2
    assert hasattr(gpt, "GPTModel") and hasattr(groups, "Topology") and hasattr(flat, "FlatParams")
    for sub in ("models", "ops", "parallel", "utils"):
        assert (PK / sub / "__init__.py").exists()
