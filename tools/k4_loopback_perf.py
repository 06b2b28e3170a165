"""K3 / K4 kernel-structure cost on ONE GPU (loopback: rank 0 of a world of 2, peer buffers local, peer flags pre-set):
fused kernel vs the plain CTA-pair GEMM at the Llama TP2 shapes.  What is left after subtracting this from the 2-GPU
numbers (tools/tp_perf.py) is waiting on the peer.  B2F_RS_DEBUG=1 (no sends / receives) and =2 (sends only) split K4."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, "/root/repo")
from b200flow.ops import _lib  # noqa: E402
from b200flow.ops import gemm as G  # noqa: E402
from b200flow.ops._lib import c_int, c_ll, ptr, stream_ptr  # noqa: E402

dev = torch.device("cuda", 0)
FUTURE = 1 << 30
T, world = 4096, 2
ts = T // world
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def ptrs(*t):
    return (ctypes.c_longlong * len(t))(*[x.data_ptr() for x in t])


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1))
    out.sort()
    return out[len(out) // 2]


epoch = [1]
for name, kind, N, K, mn in [("fwd_attn_out", "rs", 4096, 2048, 0), ("fwd_down", "rs", 4096, 5504, 0),
                             ("bwd_qkv_dgrad", "rs", 4096, 6144, 1), ("bwd_gate_up_dgrad", "rs", 4096, 11008, 1),
                             ("fwd_qkv", "ag", 6144, 4096, 0), ("fwd_gate_up", "ag", 11008, 4096, 0),
                             ("bwd_attn_out_dgrad", "ag", 2048, 4096, 1), ("bwd_down_dgrad", "ag", 5504, 4096, 1)]:
    w = (torch.randn(K, N, device=dev) if mn else torch.randn(N, K, device=dev)).to(torch.bfloat16)
    ldw = N if mn else K
    x = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
    t_g = timeit(lambda: G.gemm(x, w, b_mn_major=bool(mn)))
    if kind == "rs":
        recv = [torch.zeros(world * ts * N, device=dev, dtype=torch.bfloat16) for _ in range(world)]
        fl = [torch.full((1 << 16,), FUTURE, dtype=torch.int32, device=dev) for _ in range(world)]
        ysh = torch.empty(ts, N, device=dev, dtype=torch.bfloat16)

        def run():
            epoch[0] += 1
            _lib.call("b2f_gemm_rs_bf16", ptr(x), ptr(w), ptr(ysh), c_int(T), c_int(N), c_int(K), c_ll(K), c_ll(ldw),
                      c_ll(N), c_int(mn), ptr(None), ptr(None), c_ll(0), ptrs(*recv), ptrs(*fl), c_int(0), c_int(world),
                      ctypes.c_uint(epoch[0]), stream_ptr())
    else:
        xs = [torch.randn(ts, K, device=dev, dtype=torch.bfloat16) for _ in range(world)]
        xg = torch.empty(T, K, device=dev, dtype=torch.bfloat16)
        y = torch.empty(T, N, device=dev, dtype=torch.bfloat16)
        ready = [torch.full((64,), FUTURE, dtype=torch.int32, device=dev) for _ in range(world)]
        tile_flags = torch.zeros(4096, dtype=torch.int32, device=dev)
        ctas = int(os.environ.get("B2F_AG_CTAS", "16"))

        def run():
            epoch[0] += 1
            _lib.call("b2f_ag_gemm_bf16", ptrs(*xs), ptr(xg), ptr(w), ptr(y), c_int(T), c_int(N), c_int(K), c_ll(ldw),
                      c_ll(N), c_int(mn), ptr(None), ptrs(*ready), ptr(tile_flags), c_int(0), c_int(world), c_int(ctas),
                      ctypes.c_uint(epoch[0]), stream_ptr())
    t_f = timeit(run)
    print(json.dumps({"shape": name, "kind": kind, "N": N, "K": K, "loopback_fused_ms": round(t_f, 4),
                      "gemm_alone_ms": round(t_g, 4), "ratio": round(t_f / t_g, 3),
                      "rs_debug": os.environ.get("B2F_RS_DEBUG", "0"), "interleave": os.environ.get("B2F_RS_INTERLEAVE", "1")}),
          flush=True)
