"""Repository-named package directory.

The directory name contains hyphens, so Python cannot ``import`` it by name; the
importable package is :mod:`b200flow` (same tree: ``models/ ops/ parallel/ utils/`` plus the
launcher, engines, job-spec and checkpoint subsystems).  This directory keeps the layout the
project brief asks for and forwards to it, e.g. through ``importlib``::

    import importlib.util, sys
    spec = importlib.util.spec_from_file_location(
        "eks_ml_b200", "amazon-eks-machine-learning-with-terraform-and-kubeflow_b200/__init__.py",
        submodule_search_locations=["amazon-eks-machine-learning-with-terraform-and-kubeflow_b200"])
    mod = importlib.util.module_from_spec(spec); sys.modules["eks_ml_b200"] = mod
    spec.loader.exec_module(mod)
    from eks_ml_b200.ops import gemm
"""
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
if _root not in _sys.path:
    _sys.path.insert(0, _root)

from b200flow import *  # noqa: F401,F403,E402
import b200flow as _impl  # noqa: E402

__version__ = getattr(_impl, "__version__", "0")
