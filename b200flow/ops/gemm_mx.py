"""Block-scaled fp8 (MXFP8: e4m3 values, one UE8M0 scale per 32 consecutive K elements) on the tcgen05
`kind::mxf8f6f4.block_scale` GEMM (csrc/kernels/gemm_mx.cu, gemm_mx_core.cuh).

`quantize_mx` writes the scales directly in the tile ("atom") layout that the kernel moves to tensor memory with
`tcgen05.cp`; `MxTensor` keeps values and scales together.  The MoE expert FFNs of BASELINE.json config 5 run their
forward and dgrad GEMMs on this kernel (parallel/moe.py, csrc/comm/fused_moe.cu); wgrad stays bf16 with fp32
accumulation from the dequantised operands, since the scale blocks run along the hidden dimension, not along tokens.

Reference capability: the MoE expert FFNs of the pinned DeepSpeed 0.13.4 / Megatron-DeepSpeed a9856ce
(containers/megatron-deepspeed/Dockerfile:12-13, `--num-experts` of the gpt2_345m examples) run on cuBLAS in fp16 / bf16;
BASELINE.json config 5 asks for them in block-scaled fp8.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from . import _lib
from . import gemm as G
from ._lib import c_int, c_ll, ptr, stream_ptr


def sf_atoms(rows: int) -> int:
    return (rows + 127) // 128 + 1


def sf_bytes(rows: int, k: int) -> int:
    return (k // 128) * sf_atoms(rows) * 512


@dataclass
class MxTensor:
    q: torch.Tensor      # uint8 [batch?, rows, K]: e4m3 bits
    sf: torch.Tensor     # uint8 [batch?, K/128, atoms, 512]: UE8M0 scales in the kernel's atom layout

    @property
    def rows(self) -> int:
        return self.q.shape[-2]

    @property
    def k(self) -> int:
        return self.q.shape[-1]

    @property
    def batch(self) -> int:
        return self.q.shape[0] if self.q.dim() == 3 else 1


def quantize_mx(x: torch.Tensor) -> MxTensor:
    """bf16 [R, K] or [B, R, K] (K a multiple of 128, rows contiguous) -> MxTensor."""
    _lib.require_cuda(x)
    assert x.dtype == _lib.HALF and x.dim() in (2, 3) and x.stride(-1) == 1
    batch = x.shape[0] if x.dim() == 3 else 1
    R, K = x.shape[-2], x.shape[-1]
    assert K % 128 == 0, "MX quantisation needs K % 128 == 0"
    q = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    shape = (batch, K // 128, sf_atoms(R), 512) if x.dim() == 3 else (K // 128, sf_atoms(R), 512)
    sf = torch.zeros(shape, dtype=torch.uint8, device=x.device)
    _lib.call("b2f_quantize_mx", ptr(x), ptr(q), ptr(sf), c_int(R), c_int(K), c_ll(x.stride(-2)), c_int(batch),
              c_ll(x.stride(0) if x.dim() == 3 else 0), stream_ptr())
    return MxTensor(q, sf)


def quantize_mx_t(x: torch.Tensor) -> MxTensor:
    """MX copy of x^T in one pass: x [N, K] or [B, N, K] -> MxTensor with rows K, contracting over N
    (N a multiple of 128).  Same bits as `quantize_mx(x.transpose(-1, -2).contiguous())` without the copy."""
    _lib.require_cuda(x)
    assert x.dtype == _lib.HALF and x.dim() in (2, 3) and x.stride(-1) == 1
    batch = x.shape[0] if x.dim() == 3 else 1
    N, K = x.shape[-2], x.shape[-1]
    assert N % 128 == 0 and K % 8 == 0, "transposed MX quantisation needs N % 128 == 0 and K % 8 == 0"
    lead = (batch,) if x.dim() == 3 else ()
    q = torch.empty(lead + (K, N), dtype=torch.uint8, device=x.device)
    sf = torch.zeros(lead + (N // 128, sf_atoms(K), 512), dtype=torch.uint8, device=x.device)
    _lib.call("b2f_quantize_mx_t", ptr(x), ptr(q), ptr(sf), c_int(N), c_int(K), c_ll(x.stride(-2)), c_int(batch),
              c_ll(x.stride(0) if x.dim() == 3 else 0), stream_ptr())
    return MxTensor(q, sf)


def dequantize_mx(t: MxTensor) -> torch.Tensor:
    y = torch.empty(t.q.shape, dtype=_lib.HALF, device=t.q.device)
    R, K = t.rows, t.k
    _lib.call("b2f_dequantize_mx", ptr(t.q), ptr(t.sf), ptr(y), c_int(R), c_int(K), c_ll(K), c_int(t.batch),
              c_ll(R * K), stream_ptr())
    return y


def gemm_mx(a: MxTensor, b: MxTensor, bias=None, residual=None, aux=None, gelu: bool = False,
            out_dtype=None) -> torch.Tensor:
    """C = dequant(a) @ dequant(b)^T (+bias) (gelu; `aux` receives the pre-activation) (+residual)."""
    assert a.k == b.k and a.batch == b.batch
    M, N, K, batch = a.rows, b.rows, a.k, a.batch
    shape = (batch, M, N) if a.q.dim() == 3 else (M, N)
    out_dtype = out_dtype or _lib.HALF
    out = torch.empty(shape, dtype=out_dtype, device=a.q.device)
    flags = (G.EPI_GELU if gelu else 0) | (G.EPI_OUT_F32 if out_dtype == torch.float32 else 0)
    _lib.call("b2f_gemm_mx", ptr(a.q), ptr(a.sf), ptr(b.q), ptr(b.sf), ptr(out), c_int(M), c_int(N), c_int(K),
              c_int(batch), c_ll(N), c_ll(M * N), ptr(bias), c_ll(0), ptr(residual),
              c_ll(residual.stride(-2) if residual is not None else 0),
              c_ll(residual.stride(0) if residual is not None and residual.dim() == 3 else 0), ptr(aux),
              c_ll(aux.stride(-2) if aux is not None else 0),
              c_ll(aux.stride(0) if aux is not None and aux.dim() == 3 else 0), c_int(flags), stream_ptr())
    return out


def quantize_mx_reference(x: torch.Tensor):
    """The quantiser restated in PyTorch (fp32): returns (dequantised values, scale exponents [.., K/32])."""
    xf = x.float()
    g = xf.reshape(*xf.shape[:-1], xf.shape[-1] // 32, 32)
    amax = g.abs().amax(-1)
    t = amax / 448.0
    mant, ex = torch.frexp(t)                       # t = mant * 2^ex, mant in [0.5, 1)
    e = torch.where(mant == 0.5, ex - 1, ex)        # ceil(log2 t)
    e = torch.where(t == 0, torch.full_like(e, -126), e).clamp(-126, 126)
    scale = torch.pow(2.0, e.float())
    qv = (g / scale[..., None]).to(torch.float8_e4m3fn).float()
    return (qv * scale[..., None]).reshape(xf.shape), e
