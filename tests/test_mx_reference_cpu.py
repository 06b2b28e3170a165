"""CPU checks of the block-scaled fp8 (MXFP8) oracle and layout arithmetic that the GPU checks compare the kernels with
(`ops/gemm_mx.py::quantize_mx_reference`, `sf_atoms`, `sf_bytes`; kernel side: csrc/kernels/gemm_mx_core.cuh)."""
import math

import torch

from b200flow.ops import gemm_mx as MX


def test_reference_quantiser_properties():
    torch.manual_seed(0)
    x = torch.randn(64, 256) * torch.pow(2.0, torch.randint(-8, 8, (64, 1)).float())
    x[3, 32:64] = 0.0                                     # an all-zero group
    deq, e = MX.quantize_mx_reference(x)
    assert deq.shape == x.shape and e.shape == (64, 8)
    g = x.reshape(64, 8, 32)
    amax = g.abs().amax(-1)
    scale = torch.pow(2.0, e.float())
    nz = amax > 0
    # the scale is the smallest power of two that brings the group's largest magnitude inside e4m3's range (448)
    assert (amax[nz] / scale[nz] <= 448.0).all()
    assert (amax[nz] / (scale[nz] / 2) > 448.0).all() | (e[nz] <= -126).any()
    assert (deq.reshape(64, 8, 32)[3, 1] == 0).all()
    # e4m3 keeps 3 mantissa bits: relative error of a value near the group's maximum is below 2^-4, and every error
    # is below half an e4m3 step at the group's scale (2^-3 * 2^floor(log2 |v|) <= amax / 16)
    err = (deq - x).abs().reshape(64, 8, 32)
    assert (err <= amax[..., None] / 16 + 1e-30).all()
    # idempotent: quantising already-representable values changes nothing
    deq2, e2 = MX.quantize_mx_reference(deq)
    assert torch.equal(deq2, deq)


def test_scale_factor_layout_sizes():
    """One 512-byte atom per (128 rows, 128 K-elements) plus one spare atom row per k-block (the kernel's TMEM copy reads
    whole atoms); `sf_bytes` is what the quantiser writes and the GEMM's 1-D bulk copies read."""
    assert MX.sf_atoms(1) == 2 and MX.sf_atoms(128) == 2 and MX.sf_atoms(129) == 3
    for rows, k in [(128, 128), (200, 384), (4096, 2048), (16384, 8192)]:
        assert MX.sf_bytes(rows, k) == (k // 128) * (math.ceil(rows / 128) + 1) * 512
    # byte of (row, k) inside its atom: (row % 32) * 16 + ((row % 128) // 32) * 4 + (k % 128) // 32 - a bijection onto
    # the 512 bytes of an atom
    seen = {(r % 32) * 16 + ((r % 128) // 32) * 4 + (kk % 128) // 32 for r in range(128) for kk in range(0, 128, 32)}
    assert seen == set(range(512))
