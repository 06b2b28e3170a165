"""Single-GPU "loopback" run of the fused compute+collective kernels, for ncu.

ncu replays every kernel many times and serialises launches, so kernels that wait on a peer GPU's
kernel cannot be profiled in a real multi-rank run.  Here rank 0 of a world of 2 runs alone: the
"peer's" symmetric buffers are ordinary tensors on the same GPU and every flag the absent peer
would publish is pre-set far in the future, so each kernel executes its full code path — bulk
pulls, staged bulk sends, flag publication, tcgen05 main loop — minus the NVLink latency.
Timing under ncu is never a benchmark number; this exists for the SASS / pipeline evidence.
"""
import ctypes
import sys

import torch

sys.path.insert(0, "/root/repo")
from b200flow.ops import _lib  # noqa: E402
from b200flow.ops._lib import c_float, c_int, c_ll, ptr, stream_ptr  # noqa: E402

dev = torch.device("cuda", 0)
FUTURE = 1 << 30


def ptrs(*ts):
    return (ctypes.c_longlong * len(ts))(*[t.data_ptr() for t in ts])


def flags(n):
    return torch.full((n,), FUTURE, dtype=torch.int32, device=dev)


def k3_k4():
    T, N, K, world = 4096, 4096, 4096, 2
    ts = T // world
    xs = [torch.randn(ts, K, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    w = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
    xg = torch.empty(T, K, device=dev, dtype=torch.bfloat16)
    y = torch.empty(T, N, device=dev, dtype=torch.bfloat16)
    ready = [flags(64) for _ in range(world)]
    tile_flags = torch.zeros(4096, dtype=torch.int32, device=dev)
    _lib.call("b2f_ag_gemm_bf16", ptrs(*xs), ptr(xg), ptr(w), ptr(y), c_int(T), c_int(N), c_int(K), c_ll(K),
              c_ll(N), c_int(0), ptr(None), ptrs(*ready), ptr(tile_flags), c_int(0), c_int(world), c_int(8),
              ctypes.c_uint(1), stream_ptr())
    torch.cuda.synchronize()
    ref = torch.cat(xs).float() @ w.float().t()
    print("K3 loopback rel err", float((y.float() - ref).abs().max() / ref.abs().max()))
    x = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
    recv = [torch.zeros(world * ts * N, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    fl = [flags(1 << 16) for _ in range(world)]
    ysh = torch.empty(ts, N, device=dev, dtype=torch.bfloat16)
    _lib.call("b2f_gemm_rs_bf16", ptr(x), ptr(w), ptr(ysh), c_int(T), c_int(N), c_int(K), c_ll(K), c_ll(K),
              c_ll(N), c_int(0), ptr(None), ptr(None), c_ll(0), ptrs(*recv), ptrs(*fl), c_int(0), c_int(world),
              ctypes.c_uint(1), stream_ptr())
    torch.cuda.synchronize()
    ref = x[:ts].float() @ w.float().t()          # the absent peer's partial is zero
    print("K4 loopback rel err", float((ysh.float() - ref).abs().max() / ref.abs().max()))
    # K4' (GEMM -> all-reduce): rank 0's finished rows must appear in its own and in the peer's full output
    outs = [torch.zeros(T, N, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    done = [flags(64) for _ in range(world)]
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    bias = torch.randn(N, device=dev, dtype=torch.bfloat16)
    res = torch.randn(ts, N, device=dev, dtype=torch.bfloat16)
    for epoch in (2, 3):     # twice: the CTA counter must be back at zero for the second call
        _lib.call("b2f_gemm_ar_bf16", ptr(x), ptr(w), ptrs(*outs), c_int(T), c_int(N), c_int(K), c_ll(K), c_ll(K),
                  c_ll(N), c_int(0), ptr(bias), ptr(res), c_ll(N), ptrs(*recv), ptrs(*fl), ptrs(*done), ptr(counter),
                  c_int(0), c_int(world), ctypes.c_uint(epoch), stream_ptr())
        torch.cuda.synchronize()
        ref = ref_rows = x[:ts].float() @ w.float().t() + bias.float() + res.float()
        err = max(float((o[:ts].float() - ref_rows).abs().max() / ref.abs().max()) for o in outs)
        untouched = all(float(o[ts:].abs().max()) == 0.0 for o in outs)      # peer rows are the peer's to write
        print("K4AR loopback rel err", err if untouched and int(counter.item()) == 0 else 1.0)
        for o in outs:
            o.zero_()


def k1_k2():
    world, n = 2, 32 << 20
    g = [torch.randn(n, device=dev) for _ in range(world)]
    fl = [flags(1 << 14) for _ in range(world)]
    blocks = min(32, int(_lib.cuda_lib().b2f_comm_max_blocks()))
    _lib.call("b2f_allreduce_f32", ptrs(*g), c_ll(0), ptrs(*fl), c_int(0), c_int(world), c_ll(0), c_ll(n),
              c_float(0.5), ptr(None), c_int(blocks), ctypes.c_uint(1), stream_ptr())
    torch.cuda.synchronize()
    per = n // world
    reduced = torch.empty(per, device=dev)
    partial = torch.empty(4096, device=dev)
    sumsq = torch.zeros(1, device=dev)
    _lib.call("b2f_zero1_reduce", ptrs(*g), ptrs(*fl), c_int(0), c_int(world), c_ll(0), c_ll(per), ptr(reduced),
              ptr(partial), ptr(sumsq), c_int(blocks), ctypes.c_uint(3), stream_ptr())
    master, m, v = torch.randn(per, device=dev), torch.zeros(per, device=dev), torch.zeros(per, device=dev)
    params = [torch.empty(n, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    _lib.call("b2f_zero1_adam_gather", ptr(reduced), ptr(master), ptr(m), ptr(v), ptrs(*params), ptrs(*fl),
              c_int(0), c_int(world), c_ll(0), c_ll(per), c_float(1e-4), c_float(0.9), c_float(0.95), c_float(1e-8),
              c_float(0.1), c_int(1), ptr(None), ptr(None), c_int(blocks), ctypes.c_uint(5), stream_ptr())
    torch.cuda.synchronize()
    print("K1/K2 loopback done")


def k5():
    T, H, E, world = 16384, 2048, 8, 2
    cap = T // E * 2
    x = torch.randn(T, H, device=dev, dtype=torch.bfloat16)
    expert = torch.randint(0, E, (T,), device=dev, dtype=torch.int32)
    slot = torch.zeros(T, device=dev, dtype=torch.int32)
    for e in range(E):
        idx = (expert == e).nonzero().squeeze(-1)
        slot[idx] = torch.arange(idx.numel(), device=dev, dtype=torch.int32)
    bufs = [torch.zeros(world * (E // world) * cap * H, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    _lib.call("b2f_moe_dispatch", ptr(x), ptr(expert), ptr(slot), c_int(T), c_int(H), c_int(cap), ptr(None),
              ptrs(*bufs), c_int(world), c_int(E // world), c_int(0), stream_ptr())
    gate = torch.ones(T, device=dev)
    y = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
    _lib.call("b2f_moe_combine", ptr(None), ptrs(*bufs), c_int(world), c_int(E // world), c_int(0), ptr(expert),
              ptr(slot), ptr(gate), c_int(T), c_int(H), c_int(cap), ptr(y), stream_ptr())
    torch.cuda.synchronize()
    print("K5 loopback max |combine(dispatch(x)) - x|", float((y.float() - x.float()).abs().max()))


if __name__ == "__main__":
    which = sys.argv[1:] or ["k3k4", "k1k2", "k5"]
    if "k3k4" in which:
        k3_k4()
    if "k1k2" in which:
        k1_k2()
    if "k5" in which:
        k5()
