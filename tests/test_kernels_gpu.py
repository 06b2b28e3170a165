"""GPU numerics tests: every sm_100a kernel against a plain PyTorch fp32 reference.

The check bodies live in tools/kernel_check.py (they double as a bring-up tool that
prints error maps); each pytest case runs one group in-process and asserts that every
comparison in it passed.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(group):
    from tools import kernel_check as kc

    kc.RESULTS.clear()
    kc.GROUPS[group]()
    torch.cuda.synchronize()
    bad = [r for r in kc.RESULTS if not r.get("ok")]
    assert kc.RESULTS, f"group {group} produced no checks"
    assert not bad, f"{len(bad)} failing checks in {group}: {bad[:3]}"


@pytest.mark.parametrize("mode", ["nt", "nn", "tn", "tt"])
def test_gemm_layouts(mode):
    _run(f"gemm_{mode}")


def test_gemm_epilogues():
    _run("gemm_epi")


def test_norm_fwd_bwd():
    _run("norm")


def test_optimizer_family():
    _run("optim")


def test_elementwise_family():
    _run("elementwise")


def test_attention_forward():
    _run("attn_fwd")


def test_attention_backward():
    _run("attn_bwd")


def test_attention_dropout_forward_backward():
    _run("attn_drop")


def test_block_scaled_fp8_gemm():
    _run("mx")


def test_fused_moe_k5_single_gpu():
    _run("k5")


def test_fp8_gemm_and_linear():
    _run("fp8")


def test_moe_router_and_layer():
    _run("moe")


def test_detection_kernels():
    _run("detection")


def test_native_library_is_loaded():
    from b200flow.ops import _lib

    _lib.cuda_lib()
    maps = open("/proc/self/maps").read()
    assert "libb200flow_cuda.so" in maps


def test_ops_refuse_silent_fallback():
    """CUDA tensors must go through the native library; a missing symbol is an error."""
    from b200flow.ops import _lib

    with pytest.raises(AttributeError):
        _lib.call("b2f_this_symbol_does_not_exist")


def test_smoke_entry():
    import __graft_entry__ as g

    g.smoke()


def test_vocab_parallel_cross_entropy_kernels():
    _run("vocab_ce")


def test_gemm_cta_pair_kernel():
    _run("gemm2")


def test_kernels_are_deterministic():
    """Same inputs twice -> bit-identical outputs (no atomics in the GEMM / attention / norm / router paths;
    the wgrad accumulate is one reducing store per output element per call)."""
    from b200flow.ops import attention as A
    from b200flow.ops import gemm as G
    from b200flow.ops import norm as N
    from b200flow.parallel.moe import route_top1

    torch.manual_seed(0)
    dev = "cuda"
    a = torch.randn(1024, 512, device=dev, dtype=torch.bfloat16)
    b = torch.randn(768, 512, device=dev, dtype=torch.bfloat16)
    assert torch.equal(G.gemm(a, b), G.gemm(a, b))
    acc1 = torch.zeros(512, 512, device=dev)
    acc2 = torch.zeros(512, 512, device=dev)
    x = torch.randn(1024, 512, device=dev, dtype=torch.bfloat16)
    for acc in (acc1, acc2):
        for _ in range(3):
            G.gemm(a, x, a_mn_major=True, b_mn_major=True, out=acc, accumulate=True)
    assert torch.equal(acc1, acc2)

    qkv = torch.randn(2, 384, 3, 4, 128, device=dev, dtype=torch.bfloat16)
    outs = []
    for _ in range(2):
        t = qkv.clone().requires_grad_()
        o = A.flash_attention_qkvpacked(t, causal=True)
        o.backward(torch.ones_like(o))
        outs.append((o.detach().clone(), t.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])

    h = torch.randn(300, 1024, device=dev, dtype=torch.bfloat16)
    grads = []
    for _ in range(2):
        xi = h.clone().requires_grad_()
        g = torch.ones(1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
        y = N.norm(xi, g, None, rms=True, eps=1e-5)
        y = y[0] if isinstance(y, tuple) else y
        y.backward(torch.ones_like(y))
        grads.append((xi.grad.clone(), g.grad.clone()))
    assert torch.equal(grads[0][0], grads[1][0]) and torch.equal(grads[0][1], grads[1][1])

    logits = torch.randn(5000, 8, device=dev)
    r1, r2 = route_top1(logits, 700), route_top1(logits, 700)
    assert all(torch.equal(r1[k], r2[k]) for k in ("expert", "slot", "gate", "counts", "me", "ce"))


def test_fp16_build_of_the_kernels():
    """The -DB2F_HALF_FP16 build (IEEE fp16 storage and tensor-core operands, what `--fp16` selects) in a process of
    its own: GEMM layouts and epilogues, norms, elementwise family, optimizer, attention forward / backward."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    env = dict(os.environ, B2F_HALF="fp16")
    res = subprocess.run([sys.executable, str(root / "tools" / "kernel_check.py"), "gemm_nt", "gemm_tn", "gemm_epi", "norm",
                          "elementwise", "optim", "attn_fwd", "attn_bwd"], env=env, capture_output=True, text=True,
                         timeout=600, cwd=root)
    tail = "\n".join((res.stdout + res.stderr).splitlines()[-12:])
    assert "libb200flow_cuda_fp16" in res.stdout or res.returncode == 0, tail
    assert res.returncode == 0 and '"failed": 0' in res.stdout, tail
