"""Multi-GPU protocol checks (peer-memory arena, K1-K5, TP / PP / ZeRO-1 engines) under torchrun.

Needs >= 2 visible GPUs; the bodies live in tools/multigpu_check.py, which prints one JSON line per
check and exits non-zero when any fails."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _run(groups, nproc=2, timeout=600):
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    port = 29700 + os.getpid() % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tools" / "multigpu_check.py"), *groups]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    tail = "\n".join((res.stdout + res.stderr).splitlines()[-15:])
    assert res.returncode == 0, tail
    assert '"failed": 0' in res.stdout, tail


def test_peer_arena_and_gradient_collectives():
    _run(["arena", "allreduce", "zero1"])


def test_fused_tensor_parallel_gemms():
    _run(["fused_gemm"])


def test_tensor_parallel_without_sequence_parallel_uses_gemm_allreduce():
    _run(["tp_replicated"])


def test_tp_pp_engines_and_moe_expert_parallel():
    _run(["tp", "pp", "moe_ep"])


def test_fused_moe_k5_expert_parallel():
    """K5 (dispatch + MX-fp8 GEMM 1, MX-fp8 GEMM 2 + combine) over two GPUs against the three-phase path."""
    _run(["moe_k5"])


def test_ulysses_sequence_parallel_through_the_engine():
    """DeepSpeed-Ulysses (sequence <-> head all-to-all around the attention node) on GPUs against one GPU."""
    _run(["ulysses"])
