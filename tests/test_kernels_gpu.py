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
