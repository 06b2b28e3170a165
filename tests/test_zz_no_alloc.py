"""The allocation-free norm backward (`b2f_norm_bwd_noalloc`) against the ordinary entry point.

Who needs it: hosts that may not allocate device memory while they run — a server or trainer that
claimed the whole GPU for its own pool at start-up and gives libraries raw pointers only, a
process pinned to a fixed memory partition, a latency-critical loop that cannot risk the device
synchronisation an allocation can bring.  For them the scratch of the two-stage dgamma/dbeta
reduction sits in a fixed device array inside the library (limit: H <= 8192).

The test drives both entry points through the C ABI with ctypes — the way such a host would — on
buffers created up front, twice in a row with different shapes (the array keeps the previous
call's contents, which must not leak into the next result), checks that device memory in use does
not move across the calls, and checks that an over-limit H is refused with a message.

This file sorts last on purpose: if it fails under `-x`, every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 3e-2  # same relative-to-max tolerance the norm tests use


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


def _case(lib, rows, H, rms, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dev = "cuda"
    x = torch.randn(rows, H, device=dev, dtype=torch.bfloat16, generator=g)
    dy = torch.randn(rows, H, device=dev, dtype=torch.bfloat16, generator=g)
    dres = torch.randn(rows, H, device=dev, dtype=torch.bfloat16, generator=g)
    gamma = (1 + 0.1 * torch.randn(H, device=dev, generator=g)).to(torch.bfloat16)
    xf = x.float()
    mean = None if rms else xf.mean(-1).contiguous()
    var = xf.pow(2).mean(-1) if rms else xf.var(-1, unbiased=False)
    rstd = torch.rsqrt(var + 1e-5).contiguous()
    nparts = lib.b2f_norm_bwd_num_partials(ctypes.c_int(rows))
    scratch = torch.empty(nparts * 2 * H, device=dev, dtype=torch.float32)
    outs = []
    for _ in range(2):  # [0] = existing entry point, [1] = allocation-free one
        outs.append((torch.empty_like(x), torch.empty(H, device=dev), torch.empty(H, device=dev)))
    p = lambda t: ctypes.c_void_p(t.data_ptr() if t is not None else 0)  # noqa: E731
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc0 = lib.b2f_norm_bwd(p(dy), p(x), p(gamma), p(mean), p(rstd), p(dres), p(outs[0][0]),
                           p(outs[0][1]), p(outs[0][2]), p(scratch), ctypes.c_int(rows),
                           ctypes.c_int(H), ctypes.c_int(int(rms)), ctypes.c_int(0), stream)
    rc1 = lib.b2f_norm_bwd_noalloc(p(dy), p(x), p(gamma), p(mean), p(rstd), p(dres), p(outs[1][0]),
                                   p(outs[1][1]), p(outs[1][2]), ctypes.c_int(rows), ctypes.c_int(H),
                                   ctypes.c_int(int(rms)), ctypes.c_int(0), stream)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    assert rc0 == 0, lib.b2f_last_error()
    assert rc1 == 0, lib.b2f_last_error()
    assert _rel(outs[1][0], outs[0][0]) < TOL
    assert _rel(outs[1][1], outs[0][1]) < TOL
    if not rms:
        assert _rel(outs[1][2], outs[0][2]) < TOL
    return free_before, free_after


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    from b200flow.ops import _lib

    lib = _lib.cuda_lib()
    lib.b2f_last_error.restype = ctypes.c_char_p
    # first call on this device also resolves the scratch address (outside any capture)
    _case(lib, rows=300, H=4096, rms=True, seed=1)
    # second call: other shape and LayerNorm, so stale partials from call one would show
    _case(lib, rows=4096, H=1024, rms=False, seed=2)
    # same shape again: every kernel involved is now resident, so free memory must not move
    before, after = _case(lib, rows=4096, H=1024, rms=False, seed=4)
    assert before == after, "device memory in use changed across allocation-free calls"
    # the largest supported width
    _case(lib, rows=64, H=lib.b2f_norm_bwd_noalloc_max_h(), rms=True, seed=3)


def test_noalloc_rejects_sizes_beyond_the_limit():
    from b200flow.ops import _lib

    lib = _lib.cuda_lib()
    lib.b2f_last_error.restype = ctypes.c_char_p
    H = lib.b2f_norm_bwd_noalloc_max_h() + 8
    t = torch.zeros(8, H, device="cuda", dtype=torch.bfloat16)
    f = torch.zeros(H, device="cuda")
    p = lambda u: ctypes.c_void_p(u.data_ptr())  # noqa: E731
    rc = lib.b2f_norm_bwd_noalloc(p(t), p(t), p(t), ctypes.c_void_p(0), p(f), ctypes.c_void_p(0),
                                  p(t), p(f), p(f), ctypes.c_int(8), ctypes.c_int(H), ctypes.c_int(1),
                                  ctypes.c_int(0), ctypes.c_void_p(0))
    assert rc != 0
    assert b"8192" in lib.b2f_last_error()
