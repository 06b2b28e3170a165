"""ncu target for the round-2 kernels on one GPU: block-scaled fp8 GEMM, K5 up / down (all experts local), K3 / K4 in
loopback at the Llama TP2 shapes, attention forward with dropout.  usage: python tools/prof_r2.py [mx] [k5] [k3k4] [attn]"""
import ctypes
import sys

import torch

sys.path.insert(0, "/root/repo")
from b200flow.ops import _lib  # noqa: E402
from b200flow.ops import attention as A  # noqa: E402
from b200flow.ops import gemm_mx as MX  # noqa: E402
from b200flow.ops._lib import c_int, c_ll, ptr, stream_ptr  # noqa: E402

dev = torch.device("cuda", 0)
which = sys.argv[1:] or ["mx", "k5", "k3k4", "attn"]


def ptrs(*t):
    return (ctypes.c_longlong * len(t))(*[x.data_ptr() for x in t])


if "mx" in which:
    a = MX.quantize_mx(torch.randn(4096, 2048, device=dev, dtype=torch.bfloat16))
    b = MX.quantize_mx(torch.randn(8192, 2048, device=dev, dtype=torch.bfloat16))
    for _ in range(3):
        MX.gemm_mx(a, b)
    torch.cuda.synchronize()

if "k5" in which:
    from b200flow.parallel.moe import MoELayer

    layer = MoELayer(2048, 8192, 8, dtype=torch.bfloat16, device=dev, fp8=True)
    x = (torch.randn(16384, 2048, device=dev) * 0.5).to(torch.bfloat16).requires_grad_()
    for _ in range(2):
        layer(x).float().sum().backward()
    torch.cuda.synchronize()

if "k3k4" in which:
    FUTURE, T, world = 1 << 30, 4096, 2
    ts = T // world
    for N, K in ((6144, 4096), (4096, 5504)):
        w = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
        xs = [torch.randn(ts, K, device=dev, dtype=torch.bfloat16) for _ in range(world)]
        xg = torch.empty(T, K, device=dev, dtype=torch.bfloat16)
        y = torch.empty(T, N, device=dev, dtype=torch.bfloat16)
        ready = [torch.full((64,), FUTURE, dtype=torch.int32, device=dev) for _ in range(world)]
        tile_flags = torch.zeros(4096, dtype=torch.int32, device=dev)
        x = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
        recv = [torch.zeros(world * ts * N, device=dev, dtype=torch.bfloat16) for _ in range(world)]
        fl = [torch.full((1 << 16,), FUTURE, dtype=torch.int32, device=dev) for _ in range(world)]
        ysh = torch.empty(ts, N, device=dev, dtype=torch.bfloat16)
        for ep in (1, 2):
            _lib.call("b2f_ag_gemm_bf16", ptrs(*xs), ptr(xg), ptr(w), ptr(y), c_int(T), c_int(N), c_int(K), c_ll(K),
                      c_ll(N), c_int(0), ptr(None), ptrs(*ready), ptr(tile_flags), c_int(0), c_int(world), c_int(16),
                      ctypes.c_uint(ep), stream_ptr())
            _lib.call("b2f_gemm_rs_bf16", ptr(x), ptr(w), ptr(ysh), c_int(T), c_int(N), c_int(K), c_ll(K), c_ll(K),
                      c_ll(N), c_int(0), ptr(None), ptr(None), c_ll(0), ptrs(*recv), ptrs(*fl), c_int(0), c_int(world),
                      ctypes.c_uint(ep), stream_ptr())
    torch.cuda.synchronize()

if "attn" in which:
    qkv = torch.randn(1, 4096, 3, 32, 128, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    for _ in range(2):
        o, lse = A.attn_fwd_raw(q, k, v, True, 128 ** -0.5)
        A.attn_bwd_raw(q, k, v, o, torch.ones_like(o), lse, True, 128 ** -0.5)
    o, lse = A.attn_fwd_raw(q, k, v, True, 128 ** -0.5, None, 0.1, 7)
    A.attn_bwd_raw(q, k, v, o, torch.ones_like(o), lse, True, 128 ** -0.5, None, dropout_p=0.1, seed=7)
    torch.cuda.synchronize()
print("prof_r2 done", which)
