"""Standalone kernel bring-up checks with rich diagnostics (one group per process).

Used during development on the GPU box: each group runs in its own process under
`timeout`, so a trapped kernel cannot poison the others.  The pytest suite under
tests/ is the regression gate; this prints error maps that help debug a new
kernel blind.   usage: python tools/kernel_check.py <group> [...]
"""
from __future__ import annotations

import json
import math
import os
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.ops import elementwise as ew  # noqa: E402
from b200flow.ops import gemm as G  # noqa: E402
from b200flow.ops import norm as N  # noqa: E402
from b200flow.ops import optim as O  # noqa: E402

from b200flow.ops import _lib as _b2f_lib  # noqa: E402

DEV = "cuda"
HALF = _b2f_lib.HALF          # bf16, or fp16 when the process runs the fp16 build (B2F_HALF=fp16)
RESULTS = []


def report_rms(name, got, ref, tol):
    """Quantisation-noise comparisons (fp8 result against the unquantised product): the error of one element is a sum
    of many independent rounding errors, so single elements legitimately stray several sigma; what the format
    promises is the noise level.  Checks ||got - ref|| / ||ref|| per row (last dimension) against `tol`, and that no
    element is further off than 8 x tol x the row's rms (a systematic error - wrong scale, dropped block - is far
    beyond either)."""
    got = got.float()
    ref = ref.float().to(got.device)
    err = got - ref
    rms_ref = ref.pow(2).mean(-1, keepdim=True).sqrt().clamp_min(ref.pow(2).mean().sqrt() * 1e-3 + 1e-12)
    row_noise = (err.pow(2).mean(-1, keepdim=True).sqrt() / rms_ref).max().item()
    worst_elem = (err.abs() / rms_ref).max().item()
    ok = bool(row_noise <= tol) and bool(worst_elem <= 8 * tol) and bool(torch.isfinite(got).all())
    rec = {"name": name, "ok": ok, "row_noise": row_noise, "worst_elem": worst_elem, "tol": tol}
    RESULTS.append(rec)
    print(json.dumps(rec), flush=True)
    return ok


def report(name, got, ref, tol, floor=1e-3):
    """Per-element check  |got - ref| <= tol * (rms_row(ref) + |ref|):  the relative part bounds the rounding of the
    bf16 / fp32 result itself (1 bf16 ulp = 0.4 %), the absolute part is scaled by the rms of the element's own row
    (last dimension), not by the tensor's global maximum, so a wrong low-magnitude row or a dropped K tail cannot
    hide behind large values elsewhere.  `worst` is the largest ratio error / bound over all elements.
    `floor`: rows are held to at least this fraction of the tensor's overall rms.  Kernels whose result is a difference
    of rounded intermediates (attention backward: dS = P (dP - rowsum(O dO)) with O and P in bf16) have an error that
    scales with the tensor, not with the row - rows that attend to one or two keys have an exactly-zero gradient -
    and pass a floor of order 1."""
    got = got.float()
    ref = ref.float().to(got.device)
    err = (got - ref).abs()
    rows = ref.reshape(-1, ref.shape[-1]) if ref.dim() >= 1 and ref.numel() else ref.reshape(1, -1)
    rms = rows.pow(2).mean(-1, keepdim=True).sqrt().reshape(*ref.shape[:-1], 1) if ref.dim() >= 1 else ref.abs()
    floor = ref.pow(2).mean().sqrt() * floor + 1e-12       # an all-zero row is held to the tensor's scale
    bound = tol * (torch.maximum(rms, floor) + ref.abs())
    ratio = err / bound
    worst = ratio.max().item() if ratio.numel() else 0.0
    ok = bool(worst <= 1.0) and bool(torch.isfinite(got).all())
    denom = ref.abs().max().item() + 1e-6
    rec = {"name": name, "ok": ok, "worst": worst, "tol": tol, "max_abs": err.max().item() if err.numel() else 0.0,
           "rel_to_max": (err.max().item() / denom) if err.numel() else 0.0, "ref_absmax": denom,
           "violations": int((ratio > 1.0).sum().item())}
    if not ok and got.dim() == 2:
        bad = ratio > 1.0
        rows = bad.any(1).nonzero().flatten()
        cols = bad.any(0).nonzero().flatten()
        rec["bad_frac"] = bad.float().mean().item()
        rec["bad_rows_head"] = rows[:16].tolist()
        rec["bad_rows_count"] = int(rows.numel())
        rec["bad_cols_head"] = cols[:16].tolist()
        rec["bad_cols_count"] = int(cols.numel())
        rec["got_head"] = got[:2, :8].tolist()
        rec["ref_head"] = ref[:2, :8].tolist()
    RESULTS.append(rec)
    print(json.dumps(rec), flush=True)
    return ok


def timeit(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def group_gemm(mode):
    torch.manual_seed(0)
    shapes = [(128, 256, 64), (256, 512, 256), (384, 768, 320), (200, 264, 136), (4096, 4096, 4096),
              # odd shapes and K tails: K not a multiple of the 64-element k-block, M / N off the tile grid
              (130, 264, 72), (257, 520, 200), (129, 8, 24), (513, 1032, 328), (64, 136, 8)]
    for (M, Nn, K) in shapes:
        # an MN-major operand is a [K, MN] matrix read through TMA: its row pitch (MN elements) must be a multiple
        # of 16 bytes, so the transposed layouts take MN in multiples of 8
        if mode in ("tn", "tt"):
            M = (M + 7) // 8 * 8
        if mode in ("nn", "tn"):
            Nn = (Nn + 7) // 8 * 8
        a = torch.randn(M, K, device=DEV, dtype=HALF)
        b = torch.randn(Nn, K, device=DEV, dtype=HALF)
        ref = a.float() @ b.float().t()
        if mode == "nt":
            got = G.gemm(a, b)
        elif mode == "nn":
            got = G.gemm(a, b.t().contiguous(), b_mn_major=True)
        elif mode == "tn":
            got = G.gemm(a.t().contiguous(), b.t().contiguous(), a_mn_major=True, b_mn_major=True)
        elif mode == "tt":
            got = G.gemm(a.t().contiguous(), b, a_mn_major=True)
        torch.cuda.synchronize()
        report(f"gemm_{mode}_{M}x{Nn}x{K}", got, ref, 5e-3)      # fp32 accumulation, one bf16 rounding (0.4 %) at the end


def group_gemm_epi():
    torch.manual_seed(1)
    M, Nn, K = 512, 768, 256
    a = torch.randn(M, K, device=DEV, dtype=HALF)
    b = torch.randn(Nn, K, device=DEV, dtype=HALF) * 0.05
    bias = torch.randn(Nn, device=DEV, dtype=HALF)
    res = torch.randn(M, Nn, device=DEV, dtype=HALF)
    ref = a.float() @ b.float().t()
    report("epi_bias", G.gemm(a, b, bias=bias), ref + bias.float(), 2e-2)
    report("epi_res", G.gemm(a, b, residual=res), ref + res.float(), 2e-2)
    aux = torch.empty(M, Nn, device=DEV, dtype=HALF)
    got = G.gemm(a, b, bias=bias, gelu=True, aux=aux)
    pre = ref + bias.float()
    report("epi_gelu", got, torch.nn.functional.gelu(pre, approximate="tanh"), 2e-2)
    report("epi_aux", aux, pre, 2e-2)
    c32 = torch.ones(M, Nn, device=DEV, dtype=torch.float32)
    G.gemm(a, b, out=c32, accumulate=True)
    report("epi_f32_accum", c32, ref + 1.0, 2e-2)
    c32b = torch.empty(M, Nn, device=DEV, dtype=torch.float32)
    G.gemm(a, b, out=c32b, alpha=0.5)
    report("epi_f32_alpha", c32b, 0.5 * ref, 2e-2)
    # dgelu epilogue: C = (A B^T) * gelu'(z), z passed through the residual pointer
    z = torch.randn(M, Nn, device=DEV, dtype=HALF)
    zf = z.float().requires_grad_()
    torch.nn.functional.gelu(zf, approximate="tanh").sum().backward()
    report("epi_dgelu", G.gemm(a, b, residual=z, dgelu=True), ref * zf.grad, 2e-2)
    # fused FFN (two GEMMs forward, four backward) against fp32 autograd
    H_, F_ = 256, 1024
    x = torch.randn(384, H_, device=DEV, dtype=HALF, requires_grad=True)
    w1 = (torch.randn(F_, H_, device=DEV) * 0.05).to(HALF).requires_grad_()
    b1 = (torch.randn(F_, device=DEV) * 0.1).to(HALF).requires_grad_()
    w2 = (torch.randn(H_, F_, device=DEV) * 0.05).to(HALF).requires_grad_()
    b2 = (torch.randn(H_, device=DEV) * 0.1).to(HALF).requires_grad_()
    y = G.mlp_gelu(x, w1, b1, w2, b2)
    dy = torch.randn_like(y)
    y.backward(dy)
    fx, fw1, fb1, fw2, fb2 = (t.detach().float().requires_grad_() for t in (x, w1, b1, w2, b2))
    yr = torch.nn.functional.gelu(fx @ fw1.t() + fb1, approximate="tanh") @ fw2.t() + fb2
    yr.backward(dy.float())
    report("mlp_fwd", y, yr, 2e-2)
    for nm, got, want in (("dx", x.grad, fx.grad), ("dw1", w1.grad, fw1.grad), ("db1", b1.grad, fb1.grad),
                          ("dw2", w2.grad, fw2.grad), ("db2", b2.grad, fb2.grad)):
        report(f"mlp_{nm}", got.reshape(-1, got.shape[-1]), want.reshape(-1, want.shape[-1]), 3e-2)
    # batched
    ab = torch.randn(3, 256, 128, device=DEV, dtype=HALF)
    bb = torch.randn(3, 512, 128, device=DEV, dtype=HALF)
    report("gemm_batched", G.gemm(ab, bb).reshape(-1, 512),
           torch.bmm(ab.float(), bb.float().transpose(1, 2)).reshape(-1, 512), 2e-2)


def group_gemm_perf():
    out = {}
    for (M, Nn, K) in [(4096, 4096, 4096), (8192, 8192, 8192), (4096, 22016, 4096), (4096, 4096, 11008),
                       (4096, 12288, 4096), (4096, 32000, 4096)]:
        a = torch.randn(M, K, device=DEV, dtype=HALF)
        b = torch.randn(Nn, K, device=DEV, dtype=HALF)
        c = torch.empty(M, Nn, device=DEV, dtype=HALF)
        t_ours = timeit(lambda: G.gemm(a, b, out=c))
        t_cublas = timeit(lambda: torch.matmul(a, b.t(), out=c))
        fl = 2.0 * M * Nn * K
        rec = {"name": f"perf_{M}x{Nn}x{K}", "ours_ms": t_ours, "cublas_ms": t_cublas,
               "ours_tflops": fl / t_ours / 1e9, "cublas_tflops": fl / t_cublas / 1e9}
        # dgrad / wgrad modes
        bt = b.t().contiguous()
        t_nn = timeit(lambda: G.gemm(a, bt, out=c, b_mn_major=True))
        at = a.t().contiguous()
        t_tn = timeit(lambda: G.gemm(at, bt, out=c, a_mn_major=True, b_mn_major=True))
        rec["nn_tflops"] = fl / t_nn / 1e9
        rec["tn_tflops"] = fl / t_tn / 1e9
        print(json.dumps(rec), flush=True)
        out[rec["name"]] = rec
    RESULTS.append({"name": "gemm_perf", "ok": True, "detail": out})


def group_norm():
    torch.manual_seed(2)
    for rms in (True, False):
        for (rows, H) in [(64, 1024), (300, 4096), (4096, 4096), (17, 768), (33, 2048), (9, 8192), (70, 5120), (5, 16384)]:
            x = torch.randn(rows, H, device=DEV, dtype=HALF, requires_grad=True)
            r = torch.randn(rows, H, device=DEV, dtype=HALF, requires_grad=True)
            g = (1 + 0.1 * torch.randn(H, device=DEV)).to(HALF).requires_grad_()
            bta = None if rms else (0.1 * torch.randn(H, device=DEV)).to(HALF).requires_grad_()
            y, h = N.norm(x, g, bta, residual=r, rms=rms, eps=1e-5)
            dy = torch.randn_like(y)
            dh = torch.randn_like(h)
            (y.float() * dy.float()).sum().backward(retain_graph=True)
            (h.float() * dh.float()).sum().backward()
            xf = x.detach().float().requires_grad_()
            rf = r.detach().float().requires_grad_()
            gf = g.detach().float().requires_grad_()
            bf = bta.detach().float().requires_grad_() if bta is not None else None
            hf = (xf + rf).to(HALF).float() + 0 * (xf + rf)
            hf = xf + rf
            if rms:
                yf = hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + 1e-5) * gf
            else:
                yf = torch.nn.functional.layer_norm(hf, (H,), gf, bf, 1e-5)
            ((yf * dy.float()).sum() + (hf * dh.float()).sum()).backward()
            tag = f"{'rms' if rms else 'ln'}_{rows}x{H}"
            report(f"norm_y_{tag}", y, yf, 2e-2)
            report(f"norm_dx_{tag}", x.grad, xf.grad, 3e-2)
            report(f"norm_dres_{tag}", r.grad, rf.grad, 3e-2)
            report(f"norm_dg_{tag}", g.grad, gf.grad, 3e-2)
            if bta is not None:
                report(f"norm_db_{tag}", bta.grad, bf.grad, 3e-2)


def group_norm_perf():
    """Bandwidth of the norm kernels (bytes = algorithmic minimum) against the measured HBM peak."""
    for (rows, H, rms) in [(32768, 1024, False), (4096, 4096, True), (16384, 4096, True), (8192, 8192, True)]:
        x = torch.randn(rows, H, device=DEV, dtype=HALF, requires_grad=True)
        r = torch.randn(rows, H, device=DEV, dtype=HALF)
        g = torch.ones(H, device=DEV, dtype=HALF, requires_grad=True)
        b = None if rms else torch.zeros(H, device=DEV, dtype=HALF, requires_grad=True)
        y, h = N.norm(x, g, b, residual=r, rms=rms, eps=1e-5)
        dy = torch.randn_like(y)
        t_f = timeit(lambda: N.norm(x.detach(), g.detach(), None if b is None else b.detach(), residual=r, rms=rms, eps=1e-5))
        t_b = timeit(lambda: torch.autograd.grad(y, (x, g), dy, retain_graph=True))
        nb = rows * H * 2
        rec = {"name": f"norm_perf_{rows}x{H}_{'rms' if rms else 'ln'}", "fwd_ms": t_f, "bwd_ms": t_b,
               "fwd_GBps": 4 * nb / t_f / 1e6, "bwd_GBps": 3 * nb / t_b / 1e6}
        print(json.dumps(rec), flush=True)
        RESULTS.append({"name": rec["name"], "ok": True})


def group_vocab_ce():
    """K9 on one GPU: two vocab shards of the same logits play the two ranks."""
    import ctypes
    from b200flow.ops import _lib
    from b200flow.ops._lib import c_float, c_int, c_ll, ptr, stream_ptr
    torch.manual_seed(11)
    T, V, world = 300, 2048, 2
    Vl = V // world
    logits = (torch.randn(T, V, device=DEV) * 3).to(HALF)
    targets = torch.randint(0, V, (T,), device=DEV)
    targets[::7] = -100
    lf = logits.float().requires_grad_()
    ref = torch.nn.functional.cross_entropy(lf, targets, ignore_index=-100, reduction="sum")
    ref.backward()
    shards = [logits[:, r * Vl:(r + 1) * Vl].contiguous() for r in range(world)]
    stats = torch.zeros(world * T * 3, device=DEV)
    both = (ctypes.c_longlong * world)(*[stats.data_ptr()] * world)
    for r in range(world):
        _lib.call("b2f_vocab_ce_stats", ptr(shards[r]), ptr(targets), c_ll(T), c_int(Vl), c_ll(Vl), c_ll(r * Vl),
                  c_int(r), c_int(world), both, stream_ptr())
    losses = []
    for r in range(world):
        loss = torch.empty(T, device=DEV)
        _lib.call("b2f_vocab_ce_finish", ptr(shards[r]), ptr(targets), c_ll(T), c_int(Vl), c_ll(Vl), c_ll(r * Vl),
                  c_int(world), ptr(stats), c_ll(-100), c_float(1.0), ptr(loss), stream_ptr())
        losses.append(loss)
    torch.cuda.synchronize()
    report("vocab_ce_loss", losses[0].sum().reshape(1, 1), ref.detach().reshape(1, 1), 5e-3)
    report("vocab_ce_loss_ranks_agree", losses[0].reshape(1, -1), losses[1].reshape(1, -1), 1e-6)
    report("vocab_ce_grad", torch.cat(shards, 1), lf.grad, 2e-2)


def group_optim():
    torch.manual_seed(3)
    for n in (1000003, 4096, 7):
        p = torch.randn(n, device=DEV)
        m = torch.zeros(n, device=DEV)
        v = torch.zeros(n, device=DEV)
        pb = torch.empty(n, device=DEV, dtype=HALF)
        ref_p = torch.nn.Parameter(p.clone())
        opt = torch.optim.AdamW([ref_p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
        for step in range(1, 4):
            g = torch.randn(n, device=DEV)
            ref_p.grad = g.clone()
            opt.step()
            O.adamw_flat_(p, m, v, g, pb, lr=1e-2, beta1=0.9, beta2=0.95, eps=1e-8,
                          weight_decay=0.1, step=step)
        report(f"adamw_flat_{n}", p, ref_p.detach(), 1e-4)
        report(f"adamw_flat_bf16out_{n}", pb, ref_p.detach(), 1e-2)
    x = torch.randn(3000001, device=DEV)
    report("sumsq_f32", O.sumsq(x), (x.double() ** 2).sum().float().reshape(1), 1e-4)
    xb = x.to(HALF)
    report("sumsq_bf16", O.sumsq(xb), (xb.double() ** 2).sum().float().reshape(1), 1e-4)
    coef, nrm = O.clip_coef(O.sumsq(x), 1.0)
    report("clip_norm", nrm, x.norm().reshape(1), 1e-4)
    report("clip_coef", coef, (1.0 / (x.norm() + 1e-6)).reshape(1), 1e-4)
    y = x.clone()
    O.scale_(y, 0.5, coef)
    report("scale", y, x * 0.5 * coef, 1e-5)
    # multi-tensor
    sizes = [5, 8192, 10000, 333]
    ps = [torch.randn(s, device=DEV) for s in sizes]
    ms = [torch.zeros_like(t) for t in ps]
    vs = [torch.zeros_like(t) for t in ps]
    gs = [torch.randn(s, device=DEV).to(HALF) for s in sizes]
    pbs = [torch.empty(s, device=DEV, dtype=HALF) for s in sizes]
    refs = [torch.nn.Parameter(t.clone()) for t in ps]
    opt = torch.optim.AdamW(refs, lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)
    for r_, g_ in zip(refs, gs):
        r_.grad = g_.float()
    opt.step()
    tab = O.MultiTensorTable(ps, ms, vs, gs, pbs)
    O.adamw_multi_(tab, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=1)
    report("adamw_multi", torch.cat(ps), torch.cat([r_.detach() for r_ in refs]), 1e-4)


def group_elementwise():
    torch.manual_seed(4)
    rows, F = 300, 1024
    gu = torch.randn(rows, 2 * F, device=DEV, dtype=HALF, requires_grad=True)
    out = ew.swiglu(gu)
    d = torch.randn_like(out)
    out.backward(d)
    gf = gu.detach().float().requires_grad_()
    of = torch.nn.functional.silu(gf[:, :F]) * gf[:, F:]
    of.backward(d.float())
    report("swiglu_fwd", out, of, 2e-2)
    report("swiglu_bwd", gu.grad, gf.grad, 2e-2)
    x = torch.randn(rows, F, device=DEV, dtype=HALF, requires_grad=True)
    b = torch.randn(F, device=DEV, dtype=HALF, requires_grad=True)
    y = ew.bias_gelu(x, b)
    y.backward(d)
    xf, bf = x.detach().float().requires_grad_(), b.detach().float().requires_grad_()
    yf = torch.nn.functional.gelu(xf + bf, approximate="tanh")
    yf.backward(d.float())
    report("gelu_fwd", y, yf, 2e-2)
    report("gelu_bwd", x.grad, xf.grad, 2e-2)
    report("gelu_dbias", b.grad, bf.grad, 3e-2)
    r = torch.randn(rows, F, device=DEV, dtype=HALF)
    o0 = ew.bias_dropout_add(x.detach(), b.detach(), r, 0.0)
    report("bda_p0", o0, x.detach().float() + b.detach().float() + r.float(), 2e-2)
    xx = torch.ones(4096, 1024, device=DEV, dtype=HALF, requires_grad=True)
    o1 = ew.bias_dropout_add(xx, None, None, 0.25, seed=123)
    keep = (o1 != 0).float().mean().item()
    RESULTS.append({"name": "bda_keep_frac", "ok": abs(keep - 0.75) < 0.01, "keep": keep})
    print(json.dumps(RESULTS[-1]), flush=True)
    o1.backward(torch.ones_like(o1))
    report("bda_bwd_mask", xx.grad, o1.detach(), 1e-3)
    # rope
    tokens, nh, dh, seq = 256, 4, 128, 128
    q = torch.randn(tokens, nh, dh, device=DEV, dtype=HALF)
    cos_t, sin_t = ew.rope_tables(seq, dh, 10000.0, DEV)
    q2 = q.clone()
    ew.rope_inplace_raw(q2, cos_t, sin_t, seq, False)
    pos = torch.arange(tokens, device=DEV) % seq
    c, s = cos_t[pos][:, None, :], sin_t[pos][:, None, :]
    qf = q.float()
    x1, x2 = qf[..., : dh // 2], qf[..., dh // 2:]
    ref = torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], -1)
    report("rope_fwd", q2.reshape(tokens, -1), ref.reshape(tokens, -1), 2e-2)
    ew.rope_inplace_raw(q2, cos_t, sin_t, seq, True)
    report("rope_roundtrip", q2.reshape(tokens, -1), qf.reshape(tokens, -1), 3e-2)
    # cross entropy
    T, V = 512, 32000
    logits = (torch.randn(T, V, device=DEV) * 2).to(HALF).requires_grad_()
    tg = torch.randint(0, V, (T,), device=DEV)
    tg[:7] = -100
    lf = logits.detach().float().requires_grad_()
    ref_loss = torch.nn.functional.cross_entropy(lf, tg, ignore_index=-100)
    ref_loss.backward()
    lc = logits.detach().clone().requires_grad_()
    loss = ew.cross_entropy(lc, tg, -100)
    loss.backward()
    report("ce_loss", loss.reshape(1), ref_loss.reshape(1), 1e-3)
    report("ce_grad", lc.grad, lf.grad, 2e-2)
    # softmax
    xs = torch.randn(2, 3, 64, 64, device=DEV, dtype=HALF)
    ys = ew.scaled_masked_softmax(xs, None, 0.125, causal=True)
    mask = torch.triu(torch.ones(64, 64, device=DEV, dtype=torch.bool), 1)
    refs = torch.softmax((xs.float() * 0.125).masked_fill(mask, float("-inf")), -1)
    report("softmax_causal", ys.reshape(-1, 64), refs.reshape(-1, 64), 2e-2)
    xc = torch.randn(100003, device=DEV)
    report("cast", ew.cast_f32_bf16(xc, scale=0.5).reshape(1, -1), (xc * 0.5).reshape(1, -1), 1e-2)


def _attn_ref(q, k, v, causal, scale, kv_lens=None):
    B, S, H, D = q.shape
    HKV = k.shape[2]
    qf, kf, vf = q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)
    if HKV != H:
        kf = kf.repeat_interleave(H // HKV, 1)
        vf = vf.repeat_interleave(H // HKV, 1)
    s = qf @ kf.transpose(-1, -2) * scale
    mask = torch.zeros(B, 1, S, S, dtype=torch.bool, device=q.device)
    if causal:
        mask |= torch.triu(torch.ones(S, S, dtype=torch.bool, device=q.device), 1)
    if kv_lens is not None:
        mask |= (torch.arange(S, device=q.device)[None, :] >= kv_lens[:, None])[:, None, None, :]
    s = s.masked_fill(mask, float("-inf"))
    p = torch.softmax(s, -1)
    return (p @ vf).transpose(1, 2)


def group_attn(which):
    from b200flow.ops import attention as A
    torch.manual_seed(5)
    cfgs = [(1, 128, 2, 2, 128, False), (2, 256, 4, 4, 128, True), (1, 384, 2, 2, 64, True),
            (2, 200, 4, 2, 128, True), (1, 512, 2, 2, 128, False), (1, 1024, 4, 4, 128, True),
            (1, 896, 2, 1, 128, True, 3.0), (1, 640, 2, 2, 64, False, 3.0)]
    for cfg in cfgs:
        (B, S, H, HKV, D, causal), amp = cfg[:6], (cfg[6] if len(cfg) > 6 else 1.0)
        qkv_shape = (B, S, H + 2 * HKV, D)
        # amp > 1 makes the row max jump between key tiles (exercises the lazy O rescale)
        buf = (torch.randn(qkv_shape, device=DEV) * amp).to(HALF)
        q = buf[:, :, :H].detach().requires_grad_()
        k = buf[:, :, H:H + HKV].detach().requires_grad_()
        v = buf[:, :, H + HKV:].detach().requires_grad_()
        scale = D ** -0.5
        tag = f"B{B}S{S}H{H}kv{HKV}D{D}{'c' if causal else 'f'}" + ("amp" if amp != 1.0 else "")
        qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
        ref = _attn_ref(qf, kf, vf, causal, scale)
        if which == "fwd":
            o, lse = A.attn_fwd_raw(q.detach(), k.detach(), v.detach(), causal, scale)
            torch.cuda.synchronize()
            report(f"attn_fwd_{tag}", o.reshape(B * S, -1), ref.reshape(B * S, -1), 2e-2)
        else:
            o = A.flash_attention(q, k, v, causal, scale)
            do = torch.randn_like(o)
            o.backward(do)
            ref.backward(do.float())
            torch.cuda.synchronize()
            report(f"attn_fwd_{tag}", o.reshape(B * S, -1), ref.reshape(B * S, -1), 2e-2)
            report(f"attn_dq_{tag}", q.grad.reshape(B * S, -1), qf.grad.reshape(B * S, -1), 3e-2, floor=3.0)
            report(f"attn_dk_{tag}", k.grad.reshape(B * S, -1), kf.grad.reshape(B * S, -1), 3e-2, floor=2.0)
            report(f"attn_dv_{tag}", v.grad.reshape(B * S, -1), vf.grad.reshape(B * S, -1), 3e-2, floor=1.0)
    if which == "bwd":
        # key-padding lengths + packed qkv path
        B, S, H, D = 2, 256, 2, 128
        qkv = torch.randn(B, S, 3, H, D, device=DEV, dtype=HALF, requires_grad=True)
        lens = torch.tensor([256, 130], device=DEV, dtype=torch.int32)
        o = A.flash_attention_qkvpacked(qkv, False, D ** -0.5, lens)
        do = torch.randn_like(o)
        o.backward(do)
        qf = qkv.detach().float().requires_grad_()
        ref = _attn_ref(qf[:, :, 0], qf[:, :, 1], qf[:, :, 2], False, D ** -0.5, lens)
        ref.backward(do.float())
        report("attn_packed_fwd", o.reshape(B * S, -1), ref.reshape(B * S, -1), 2e-2)
        report("attn_packed_dqkv", qkv.grad.reshape(B * S, -1), qf.grad.reshape(B * S, -1), 3e-2, floor=2.0)


def group_attn_perf():
    from b200flow.ops import attention as A
    for (B, S, H, D, causal) in [(1, 4096, 32, 128, True), (8, 512, 16, 64, False)]:
        qkv = torch.randn(B, S, 3, H, D, device=DEV, dtype=HALF)
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        scale = D ** -0.5
        o, lse = A.attn_fwd_raw(q, k, v, causal, scale)
        do = torch.randn_like(o)
        t_f = timeit(lambda: A.attn_fwd_raw(q, k, v, causal, scale))
        t_b = timeit(lambda: A.attn_bwd_raw(q, k, v, o, do, lse, causal, scale))
        fl = 4.0 * B * H * S * S * D * (0.5 if causal else 1.0)
        rec = {"name": f"attn_perf_B{B}S{S}H{H}D{D}", "fwd_ms": t_f, "bwd_ms": t_b,
               "fwd_tflops": fl / t_f / 1e9, "bwd_tflops": 2.5 * fl / t_b / 1e9}
        try:
            qt, kt, vt = (t.transpose(1, 2) for t in (q, k, v))
            t_ref = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=causal))
            rec["sdpa_fwd_ms"] = t_ref
            rec["sdpa_fwd_tflops"] = fl / t_ref / 1e9
        except Exception as e:  # noqa: BLE001
            rec["sdpa_error"] = repr(e)[:200]
        # backward of the libraries this kernel replaces: SDPA (cuDNN / flash backend torch picks) and flash_attn 2.8
        try:
            qs, ks, vs = (t.transpose(1, 2).detach().clone().requires_grad_() for t in (q, k, v))
            os_ = torch.nn.functional.scaled_dot_product_attention(qs, ks, vs, is_causal=causal)
            gs = torch.randn_like(os_)
            t_rb = timeit(lambda: torch.autograd.grad(os_, (qs, ks, vs), gs, retain_graph=True))
            rec["sdpa_bwd_ms"] = t_rb
            rec["sdpa_bwd_tflops"] = 2.5 * fl / t_rb / 1e9
        except Exception as e:  # noqa: BLE001
            rec["sdpa_bwd_error"] = repr(e)[:200]
        try:
            from flash_attn import flash_attn_func

            qf, kf, vf = (t.detach().clone().requires_grad_() for t in (q, k, v))
            t_ff = timeit(lambda: flash_attn_func(qf, kf, vf, causal=causal))
            of = flash_attn_func(qf, kf, vf, causal=causal)
            gf = torch.randn_like(of)
            t_fb = timeit(lambda: torch.autograd.grad(of, (qf, kf, vf), gf, retain_graph=True))
            rec.update(flash_attn_fwd_ms=t_ff, flash_attn_fwd_tflops=fl / t_ff / 1e9, flash_attn_bwd_ms=t_fb,
                       flash_attn_bwd_tflops=2.5 * fl / t_fb / 1e9)
        except Exception as e:  # noqa: BLE001
            rec["flash_attn_error"] = repr(e)[:200]
        print(json.dumps(rec), flush=True)
        RESULTS.append({"name": rec["name"], "ok": True})


def group_detection():
    import torchvision
    from b200flow.ops import detection as Dt
    torch.manual_seed(6)
    # --- conv epilogue of the folded-BN backbone: y = relu(y + bias (+ residual)) in place, with gradients
    for (n, c, h, w), with_res, relu in [((2, 64, 50, 84), False, True), ((3, 256, 25, 42), True, True),
                                         ((2, 128, 13, 21), False, False)]:
        x0 = torch.randn(n, c, h, w, device=DEV).to(HALF).contiguous(memory_format=torch.channels_last)
        bias = torch.randn(c, device=DEV).to(HALF)
        res = torch.randn(n, c, h, w, device=DEV).to(HALF).contiguous(memory_format=torch.channels_last) if with_res else None
        xa = x0.clone().requires_grad_()
        ra = res.clone().requires_grad_() if with_res else None
        ya = Dt.bias_act_(xa * 1.0, bias, ra, relu)                   # * 1.0: a non-leaf the kernel may overwrite
        xb = x0.float().requires_grad_()
        rb = res.float().requires_grad_() if with_res else None
        yb = xb + bias.float().view(1, -1, 1, 1) + (rb if with_res else 0.0)
        yb = torch.relu(yb) if relu else yb
        g = torch.randn_like(yb)
        ya.backward(g.to(HALF))
        yb.backward(g)
        tag = f"{c}x{h}x{w}_res{int(with_res)}_relu{int(relu)}"
        report(f"bias_act_fwd_{tag}", ya.permute(0, 2, 3, 1).reshape(-1, c), yb.permute(0, 2, 3, 1).reshape(-1, c), 1e-2)
        report(f"bias_act_dx_{tag}", xa.grad.permute(0, 2, 3, 1).reshape(-1, c), xb.grad.permute(0, 2, 3, 1).reshape(-1, c), 1e-2)
        if with_res:
            report(f"bias_act_dres_{tag}", ra.grad.permute(0, 2, 3, 1).reshape(-1, c), rb.grad.permute(0, 2, 3, 1).reshape(-1, c), 1e-2)
    # column sums (bias gradients) against fp32
    from b200flow.ops import elementwise as Ew
    for rows, cols in [(32768, 1024), (4096, 3072), (1000, 4096), (7, 8), (5000, 264)]:
        x = torch.randn(rows, cols, device=DEV).to(HALF)
        report(f"colsum_{rows}x{cols}", Ew.colsum(x)[None], x.float().sum(0)[None], 1e-2)
    x = torch.randn(32768, 1024, device=DEV).to(HALF)
    ms_a, ms_b = timeit(lambda: Ew.colsum(x)), timeit(lambda: x.sum(0))
    print(json.dumps({"name": "colsum_perf_32768x1024", "ms": ms_a, "torch_sum_ms": ms_b, "gbs": x.numel() * 2 / ms_a / 1e6}), flush=True)
    # trainable bias through conv_bias_act: output and all three gradients against nn.Conv2d + relu in fp32
    conv = torch.nn.Conv2d(64, 256, 3, padding=1).to(DEV).to(HALF).to(memory_format=torch.channels_last)
    ref = torch.nn.Conv2d(64, 256, 3, padding=1).to(DEV)
    ref.load_state_dict({k: v.float() for k, v in conv.state_dict().items()})
    x0 = torch.randn(2, 64, 40, 56, device=DEV).to(HALF).contiguous(memory_format=torch.channels_last)
    xa, xb = x0.clone().requires_grad_(), x0.float().requires_grad_()
    # no ReLU here: a bf16 and an fp32 convolution disagree on the sign of outputs near zero, which flips mask bits and
    # shows up as noise in every gradient; the ReLU backward is checked above on identical inputs
    report("conv_bias_act_relu_fwd", Dt.conv_bias_act(x0, conv, relu=True).permute(0, 2, 3, 1).reshape(-1, 256),
           torch.relu(ref(x0.float())).permute(0, 2, 3, 1).reshape(-1, 256), 2e-2)
    ya, yb = Dt.conv_bias_act(xa, conv, relu=False), ref(xb)
    g = torch.randn_like(yb)
    ya.backward(g.to(HALF))
    yb.backward(g)
    report("conv_bias_act_fwd", ya.permute(0, 2, 3, 1).reshape(-1, 256), yb.permute(0, 2, 3, 1).reshape(-1, 256), 2e-2)
    report("conv_bias_act_dbias", conv.bias.grad[None], ref.bias.grad[None], 2e-2)
    report("conv_bias_act_dw", conv.weight.grad.reshape(256, -1), ref.weight.grad.reshape(256, -1), 3e-2)
    report("conv_bias_act_dx", xa.grad.permute(0, 2, 3, 1).reshape(-1, 64), xb.grad.permute(0, 2, 3, 1).reshape(-1, 64), 3e-2)
    # --- K12 RoIAlign fwd/bwd vs torchvision CPU (fp32), multilevel
    C = 64
    shapes = [(2, 200, 336), (2, 100, 168), (2, 50, 84), (2, 25, 42)]
    scales = [0.25, 0.125, 0.0625, 0.03125]
    feats = [torch.randn(n, h, w, C, device=DEV).to(HALF).requires_grad_() for n, h, w in shapes]
    R = 300
    xy = torch.rand(R, 2, device=DEV) * torch.tensor([1000.0, 600.0], device=DEV)
    wh = torch.rand(R, 2, device=DEV) * 500 + 8
    rois = torch.cat([torch.randint(0, 2, (R, 1), device=DEV).float(), xy, xy + wh], 1)
    for P in (7, 14):
        out = Dt.roi_align(feats, scales, rois, P)
        dout = torch.randn_like(out)
        for f in feats:
            f.grad = None
        out.backward(dout)
        lv = Dt.fpn_level_of(rois.cpu())
        ref = torch.zeros(R, P, P, C)
        cf = [f.detach().float().cpu().permute(0, 3, 1, 2).requires_grad_() for f in feats]
        parts = []
        for i in range(4):
            idx = (lv == i).nonzero().flatten()
            if idx.numel():
                o = torchvision.ops.roi_align(cf[i], rois.cpu()[idx], P, scales[i], 2, True)
                ref[idx] = o.permute(0, 2, 3, 1)
                parts.append((o * dout.float().cpu()[idx].permute(0, 3, 1, 2)).sum())
        sum(parts).backward()
        report(f"roi_align_fwd_P{P}", out.reshape(R, -1), ref.reshape(R, -1), 2e-2)
        for i in range(4):
            if cf[i].grad is not None:
                report(f"roi_align_bwd_P{P}_L{i}", feats[i].grad.permute(0, 3, 1, 2).reshape(2, -1),
                       cf[i].grad.reshape(2, -1), 3e-2)
    # --- K13 decode + NMS
    n = 5000
    anchors = torch.rand(n, 2, device=DEV) * 800
    anchors = torch.cat([anchors, anchors + torch.rand(n, 2, device=DEV) * 200 + 4], 1)
    deltas = torch.randn(n, 4, device=DEV) * 0.3
    got = Dt.box_decode(anchors, deltas, (10.0, 10.0, 5.0, 5.0), 1344.0, 800.0)
    report("box_decode", got, Dt.box_decode(anchors.cpu(), deltas.cpu(), (10.0, 10.0, 5.0, 5.0), 1344.0, 800.0), 1e-4)
    segs = [2000, 1333, 64, 1, 700]
    boxes = []
    for m in segs:
        c = torch.rand(m, 2, device=DEV) * 600
        boxes.append(torch.cat([c, c + torch.rand(m, 2, device=DEV) * 150 + 10], 1))
    boxes = torch.cat(boxes)
    off = torch.tensor([0] + segs, device=DEV).cumsum(0).int()
    keep = Dt.nms_batched(boxes, off, 0.7)
    ref_keep = Dt.nms_batched(boxes.cpu(), off.cpu(), 0.7)
    report("nms_batched", keep.float().reshape(1, -1), ref_keep.float().reshape(1, -1), 1e-6)
    # --- K14 matcher / IoU
    gts = boxes[:40].clone()
    gt_off = torch.tensor([0, 25, 40], device=DEV).int()
    pb = torch.randint(0, 2, (3000,), device=DEV).int()
    bi, bj = Dt.match_proposals(boxes[:3000], pb, gts, gt_off)
    rbi, rbj = Dt.match_proposals(boxes[:3000].cpu(), pb.cpu(), gts.cpu(), gt_off.cpu())
    report("match_iou", bi.reshape(1, -1), rbi.reshape(1, -1), 1e-4)
    agree = ((bj.cpu() == rbj) | (rbi < 1e-6)).float().mean().item()
    RESULTS.append({"name": "match_idx", "ok": agree > 0.999, "agree": agree}); print(json.dumps(RESULTS[-1]), flush=True)
    report("pairwise_iou", Dt.pairwise_iou(boxes[:500], gts), torchvision.ops.box_iou(boxes[:500].cpu(), gts.cpu()), 1e-4)
    # --- K15 mask crop / paste
    H, W = 200, 300
    masks = (torch.rand(6, H, W, device=DEV) > 0.5).to(torch.uint8)
    masks[:, 50:150, 80:220] = 1
    rois4 = torch.tensor([[60.0, 40, 250, 170], [0, 0, 299, 199], [100, 90, 130, 120], [10.5, 20.25, 80.75, 99.5]], device=DEV)
    gi = torch.tensor([0, 3, 5, 2], device=DEV).int()
    got = Dt.mask_crop(masks, rois4, gi, 28)
    ref = Dt.mask_crop_ref(masks.cpu(), rois4.cpu(), gi.cpu(), 28)
    frac = (got.cpu() == ref).float().mean().item()
    RESULTS.append({"name": "mask_crop", "ok": frac > 0.99, "agree": frac}); print(json.dumps(RESULTS[-1]), flush=True)
    probs = torch.rand(4, 28, 28, device=DEV)
    got = Dt.mask_paste(probs, rois4, H, W)
    ref = Dt.mask_paste_ref(probs.cpu(), rois4.cpu(), H, W)
    frac = (got.cpu() == ref).float().mean().item()
    RESULTS.append({"name": "mask_paste", "ok": frac > 0.995, "agree": frac}); print(json.dumps(RESULTS[-1]), flush=True)


    # K16 polygon rasteriser against the pixel-centre even-odd oracle
    from b200flow.ops import detection as Dm
    g = torch.Generator().manual_seed(5)
    objs = []
    for o in range(6):
        parts = []
        for _ in range(1 + o % 2):
            n = 3 + int(torch.randint(0, 6, (1,), generator=g))
            cx, cy = (torch.rand(2, generator=g) * 100 + 30).tolist()
            ang = torch.sort(torch.rand(n, generator=g) * 6.2831853)[0]
            rad = torch.rand(n, generator=g) * 25 + 8
            pts = torch.stack([cx + rad * torch.cos(ang), cy + rad * torch.sin(ang)], 1).reshape(-1)
            parts.append(pts.tolist())
        objs.append(parts)
    got = Dm.poly_to_mask(objs, 160, 200, device=DEV)
    want = Dm.poly_to_mask_ref(objs, 160, 200)
    diff = int((got.cpu() != want).sum())       # fp contraction may move a pixel lying exactly on an edge
    RESULTS.append({"name": "poly_to_mask", "ok": diff <= 4 and int(want.sum()) > 1000, "mismatched_pixels": diff,
                    "set_pixels": int(want.sum())})
    print(json.dumps(RESULTS[-1]), flush=True)

def group_maskrcnn_step():
    from b200flow.data.synthetic import coco_like
    from b200flow.models.maskrcnn import MaskRCNN, MaskRCNNConfig, paste_masks
    from b200flow.optim import FlatMomentumSGD
    from b200flow.utils.flat import FlatParams
    torch.manual_seed(0)
    model = MaskRCNN(MaskRCNNConfig()).to(DEV).to(HALF).to(memory_format=torch.channels_last)
    flat = FlatParams(model)
    opt = FlatMomentumSGD(flat, lr=0.005)
    data = list(coco_like(2, 800, 1344, seed=3))
    imgs = torch.stack([s["image"] for s in data]).to(DEV).to(HALF)
    tg = [{k: s[k].to(DEV) for k in ("boxes", "labels", "masks")} for s in data]
    model.train()
    ls = []
    for _ in range(4):
        opt.zero_grad()
        l = model(imgs, tg)
        l["total"].backward()
        flat.collect_autograd_grads()
        opt.step()
        ls.append(float(l["total"]))
    RESULTS.append({"name": "maskrcnn_r50fpn_train", "ok": ls[-1] < ls[0] and all(x == x for x in ls), "losses": ls})
    print(json.dumps(RESULTS[-1]), flush=True)
    def step():
        opt.zero_grad(); l = model(imgs, tg); l["total"].backward(); flat.collect_autograd_grads(); opt.step()
    t = timeit(step, iters=5, warmup=2)
    model.eval()
    with torch.no_grad():
        det = model(imgs)
    pm = paste_masks(det[0], 800, 1344)
    RESULTS.append({"name": "maskrcnn_r50fpn_perf", "ok": True, "ms_per_step_bs2": t, "images_per_s": 2 / (t / 1e3),
                    "dets": int(det[0]["boxes"].shape[0]), "pasted": list(pm.shape)})
    print(json.dumps(RESULTS[-1]), flush=True)


def group_moe():
    from b200flow.parallel.moe import MoELayer, route_top1
    torch.manual_seed(8)
    T, E, C = 3000, 8, 384
    logits = torch.randn(T, E, device=DEV)
    r = route_top1(logits, C)
    rc = route_top1(logits.cpu(), C)
    for k in ("expert", "slot", "counts"):
        RESULTS.append({"name": f"router_{k}", "ok": bool(torch.equal(r[k].cpu(), rc[k]))}); print(json.dumps(RESULTS[-1]), flush=True)
    report("router_gate", r["gate"].reshape(1, -1), rc["gate"].reshape(1, -1), 1e-5)
    report("router_me", r["me"].reshape(1, -1), rc["me"].reshape(1, -1), 1e-4)
    report("router_ce", r["ce"].reshape(1, -1), rc["ce"].reshape(1, -1), 1e-6)
    layer = MoELayer(256, 512, 8, dtype=HALF, device=DEV)
    x = (torch.randn(1024, 256, device=DEV) * 0.5).to(HALF).requires_grad_()
    y = layer(x)
    (y.float().square().sum() + layer.last_aux).backward()
    ref_layer = MoELayer(256, 512, 8, dtype=torch.float32)
    ref_layer.load_state_dict({k: v.float().cpu() for k, v in layer.state_dict().items()})
    xr = x.detach().float().cpu().requires_grad_()
    yr = ref_layer(xr)
    (yr.square().sum() + ref_layer.last_aux).backward()
    same_route = float((layer.last_counts.cpu() == ref_layer.last_counts).float().mean())
    RESULTS.append({"name": "moe_route_agree", "ok": same_route > 0.5, "agree": same_route}); print(json.dumps(RESULTS[-1]), flush=True)
    report("moe_fwd", y, yr, 4e-2)
    report("moe_dx", x.grad, xr.grad, 6e-2)
    report("moe_dw1", layer.w1.grad.reshape(8, -1), ref_layer.w1.grad.reshape(8, -1), 6e-2)
    report("moe_dwg", layer.wg.grad[:8], ref_layer.wg.grad[:8], 8e-2)


def group_attn_drop():
    """Attention dropout inside the flash kernels: forward and all three gradients against fp32 attention with the
    same mask (the hash restated in PyTorch), causal and padded, both head dims, GQA; same seed twice = same result."""
    from b200flow.ops import attention as A
    torch.manual_seed(4)
    for (B, S, H, HKV, D, causal, pd) in [(2, 512, 4, 4, 64, True, 0.1), (1, 896, 4, 2, 128, True, 0.1),
                                          (2, 384, 2, 2, 128, False, 0.25)]:
        tag = f"B{B}S{S}H{H}kv{HKV}D{D}{'c' if causal else 'f'}p{pd}"
        q = torch.randn(B, S, H, D, device=DEV, dtype=HALF).requires_grad_()
        k = torch.randn(B, S, HKV, D, device=DEV, dtype=HALF).requires_grad_()
        v = torch.randn(B, S, HKV, D, device=DEV, dtype=HALF).requires_grad_()
        kv_lens = torch.tensor([S - 37 * (i % 2) for i in range(B)], dtype=torch.int32, device=DEV)
        seed = 12345 + S
        o = A.flash_attention(q, k, v, causal, None, kv_lens, dropout_p=pd, seed=seed)
        do = torch.randn_like(o)
        o.backward(do)
        o2 = A.flash_attention(q.detach(), k.detach(), v.detach(), causal, None, kv_lens, dropout_p=pd, seed=seed)
        RESULTS.append({"name": f"attn_drop_repeat_{tag}", "ok": bool(torch.equal(o, o2))}); print(json.dumps(RESULTS[-1]), flush=True)
        keep, rinv = A.dropout_keep_mask(B, H, S, pd, seed, DEV)
        frac = 1.0 - keep.float().mean().item()
        RESULTS.append({"name": f"attn_drop_rate_{tag}", "ok": abs(frac - round(pd * 256) / 256) < 0.01, "dropped": frac})
        print(json.dumps(RESULTS[-1]), flush=True)
        qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
        G_ = H // HKV
        s_ = torch.einsum("bqhd,bkhd->bhqk", qf, kf.repeat_interleave(G_, 2)) / math.sqrt(D)
        mask = torch.zeros(B, 1, S, S, dtype=torch.bool, device=DEV)
        if causal:
            mask = mask | torch.triu(torch.ones(S, S, dtype=torch.bool, device=DEV), 1)
        mask = mask | (torch.arange(S, device=DEV)[None, :] >= kv_lens[:, None])[:, None, None, :]
        pr = torch.softmax(s_.masked_fill(mask, float("-inf")), -1) * keep * rinv
        ref = torch.einsum("bhqk,bkhd->bqhd", pr, vf.repeat_interleave(G_, 2))
        ref.backward(do.float())
        report(f"attn_drop_fwd_{tag}", o.reshape(B * S, -1), ref.reshape(B * S, -1), 2e-2, floor=0.5)
        report(f"attn_drop_dq_{tag}", q.grad.reshape(B * S, -1), qf.grad.reshape(B * S, -1), 3e-2, floor=4.0)
        report(f"attn_drop_dk_{tag}", k.grad.reshape(B * S, -1), kf.grad.reshape(B * S, -1), 3e-2, floor=2.0)
        report(f"attn_drop_dv_{tag}", v.grad.reshape(B * S, -1), vf.grad.reshape(B * S, -1), 3e-2, floor=1.0)
    # cost of the mask: B1 S4096 H32 D128 causal, forward + backward
    q = torch.randn(1, 4096, 32, 128, device=DEV, dtype=HALF).requires_grad_()
    k, v = (torch.randn(1, 4096, 32, 128, device=DEV, dtype=HALF).requires_grad_() for _ in range(2))
    def run(pd):
        o = A.flash_attention(q, k, v, True, None, None, dropout_p=pd, seed=7)
        o.backward(o)
    t0, t1 = timeit(lambda: run(0.0)), timeit(lambda: run(0.1))
    print(json.dumps({"name": "attn_drop_cost_B1S4096H32D128", "fwd_bwd_ms_no_dropout": t0, "fwd_bwd_ms_dropout": t1}), flush=True)


def group_k5():
    """K5 on one GPU (all experts local): the fused dispatch + block-scaled GEMM / GEMM + combine kernels against the fp32
    layer, forward and backward, with an uneven routing (some experts overflow, one nearly empty) and twice in a row
    (buffer parity, counters)."""
    from b200flow.parallel.moe import MoELayer
    torch.manual_seed(8)
    H, F, E, T = 512, 1024, 8, 4096
    layer = MoELayer(H, F, E, dtype=HALF, device=DEV, fp8=True, capacity_factor=1.0)
    with torch.no_grad():
        layer.b1.normal_(0, 0.05)
        layer.b2.normal_(0, 0.05)
        layer.wg[1] *= 3.0          # expert 1 attracts more tokens than its capacity, others run short
    assert layer._use_k5(torch.empty(T, H, device=DEV), layer.capacity(T)), "K5 path not taken"
    ref_layer = MoELayer(H, F, E, dtype=torch.float32, capacity_factor=1.0)
    ref_layer.load_state_dict({k: v.float().cpu() for k, v in layer.state_dict().items()})
    for rep in range(2):
        x = (torch.randn(T, H, device=DEV) * 0.5).to(HALF).requires_grad_()
        y = layer(x)
        (y.float().square().sum() + layer.last_aux).backward()
        torch.cuda.synchronize()
        xr = x.detach().float().cpu().requires_grad_()
        yr = ref_layer(xr)
        (yr.square().sum() + ref_layer.last_aux).backward()
        agree = float((layer.last_counts.cpu() == ref_layer.last_counts).float().mean())
        RESULTS.append({"name": f"k5_route_agree_r{rep}", "ok": agree > 0.99, "agree": agree,
                        "counts": layer.last_counts.tolist()})
        print(json.dumps(RESULTS[-1]), flush=True)
        report_rms(f"k5_fwd_r{rep}", y, yr, 6e-2)
        report_rms(f"k5_dx_r{rep}", x.grad, xr.grad, 8e-2)
        report_rms(f"k5_dw1_r{rep}", layer.w1.grad.reshape(E, -1), ref_layer.w1.grad.reshape(E, -1), 8e-2)
        report_rms(f"k5_dw2_r{rep}", layer.w2.grad.reshape(E, -1), ref_layer.w2.grad.reshape(E, -1), 8e-2)
        report_rms(f"k5_db1_r{rep}", layer.b1.grad.reshape(1, -1), ref_layer.b1.grad.reshape(1, -1), 8e-2)
        report_rms(f"k5_db2_r{rep}", layer.b2.grad.reshape(1, -1), ref_layer.b2.grad.reshape(1, -1), 8e-2)
        report_rms(f"k5_dwg_r{rep}", layer.wg.grad[:E], ref_layer.wg.grad[:E], 1e-1)
        layer.zero_grad(); ref_layer.zero_grad()
    # the size of BASELINE config 5 on one GPU: 8 experts x 2048 tokens each
    big = MoELayer(2048, 8192, 8, dtype=HALF, device=DEV, fp8=True)
    xb = (torch.randn(16384, 2048, device=DEV) * 0.5).to(HALF).requires_grad_()
    def step():
        big(xb).float().sum().backward()
    ms = timeit(step, iters=5, warmup=2)
    flops = 3 * 2 * 2 * 16384 * 2048 * 8192
    print(json.dumps({"name": "k5_layer_fwd_bwd_16384x2048x8192", "ms": ms, "model_tflops": flops / ms / 1e9}), flush=True)
    os.environ["B2F_FUSED_MOE"] = "0"
    try:
        ms0 = timeit(step, iters=5, warmup=2)
    finally:
        os.environ["B2F_FUSED_MOE"] = "1"
    print(json.dumps({"name": "unfused_rowscaled_layer_fwd_bwd_16384x2048x8192", "ms": ms0}), flush=True)


def group_fp8():
    from b200flow.ops import gemm_fp8 as F8
    torch.manual_seed(9)
    for (M, Nn, K) in [(256, 512, 256), (1000, 2048, 1024), (4096, 8192, 2048)]:
        a = (torch.randn(M, K, device=DEV) * torch.rand(M, 1, device=DEV) * 3).to(HALF)
        b = (torch.randn(Nn, K, device=DEV) * 0.05).to(HALF)
        aq, sa = F8.quantize_rows(a)
        bq, sb = F8.quantize_rows(b)
        deq = aq.view(torch.float8_e4m3fn).float() * sa[:, None]
        report(f"fp8_quant_{M}x{K}", deq, a.float(), 0.07)
        exact = (aq.view(torch.float8_e4m3fn).float() @ bq.view(torch.float8_e4m3fn).float().t()) * sa[:, None] * sb[None, :]
        got = F8.gemm_fp8(aq, sa, bq, sb)
        report(f"fp8_gemm_exact_{M}x{Nn}x{K}", got, exact, 1e-2)       # vs the same quantised operands
        report_rms(f"fp8_gemm_vs_bf16_{M}x{Nn}x{K}", got, a.float() @ b.float().t(), 6e-2)
    x = (torch.randn(2048, 2048, device=DEV)).to(HALF).requires_grad_()
    w = (torch.randn(8192, 2048, device=DEV) * 0.02).to(HALF).requires_grad_()
    bias = torch.zeros(8192, device=DEV, dtype=HALF, requires_grad=True)
    y = F8.linear_gelu_fp8(x, w, bias)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, wf = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    yf = torch.nn.functional.gelu(xf @ wf.t(), approximate="tanh")
    yf.backward(dy.float())
    report_rms("fp8_linear_gelu_fwd", y, yf, 6e-2)
    report_rms("fp8_linear_gelu_dx", x.grad, xf.grad, 8e-2)
    report_rms("fp8_linear_gelu_dw", w.grad, wf.grad, 6e-2)
    a = torch.randn(4096, 4096, device=DEV, dtype=HALF)
    b = torch.randn(8192, 4096, device=DEV, dtype=HALF)
    aq, sa = F8.quantize_rows(a); bq, sb = F8.quantize_rows(b)
    t8 = timeit(lambda: F8.gemm_fp8(aq, sa, bq, sb))
    t16 = timeit(lambda: G.gemm(a, b))
    fl = 2.0 * 4096 * 8192 * 4096
    RESULTS.append({"name": "fp8_perf", "ok": True, "fp8_tflops": fl / t8 / 1e9, "bf16_tflops": fl / t16 / 1e9})
    print(json.dumps(RESULTS[-1]), flush=True)


def group_gemm2():
    """2-CTA kernel: all layouts + epilogues vs fp32, then perf vs the 1-CTA kernel and cuBLAS."""
    G.GEMM_IMPL = "2cta"
    for mode in ("nt", "nn", "tn", "tt"):
        group_gemm(mode)
    group_gemm_epi()
    out = {}
    for (M, Nn, K) in [(4096, 4096, 4096), (8192, 8192, 8192), (4096, 22016, 4096), (4096, 4096, 11008),
                       (4096, 12288, 4096), (4096, 32000, 4096)]:
        a = torch.randn(M, K, device=DEV, dtype=HALF)
        b = torch.randn(Nn, K, device=DEV, dtype=HALF)
        c = torch.empty(M, Nn, device=DEV, dtype=HALF)
        fl = 2.0 * M * Nn * K
        G.GEMM_IMPL = "2cta"
        t2 = timeit(lambda: G.gemm(a, b, out=c))
        bt, at = b.t().contiguous(), a.t().contiguous()
        t2nn = timeit(lambda: G.gemm(a, bt, out=c, b_mn_major=True))
        t2tn = timeit(lambda: G.gemm(at, bt, out=c, a_mn_major=True, b_mn_major=True))
        G.GEMM_IMPL = "1cta"
        t1 = timeit(lambda: G.gemm(a, b, out=c))
        tc = timeit(lambda: torch.matmul(a, b.t(), out=c))
        rec = {"name": f"perf2_{M}x{Nn}x{K}", "cta2_tflops": fl / t2 / 1e9, "cta2_nn": fl / t2nn / 1e9,
               "cta2_tn": fl / t2tn / 1e9, "cta1_tflops": fl / t1 / 1e9, "cublas_tflops": fl / tc / 1e9}
        print(json.dumps(rec), flush=True)
    G.GEMM_IMPL = "2cta"
    RESULTS.append({"name": "gemm2_perf", "ok": True})


def group_mx():
    """Block-scaled fp8 GEMM (kind::mxf8f6f4, scale factors in TMEM) against fp32 products of the dequantised operands.
    The structured cases separate the parts of the scale-factor layout: uniform magnitudes need no layout at all,
    then only A rows, only B rows (output columns), only K groups vary."""
    from b200flow.ops import gemm_mx as MX

    torch.manual_seed(3)

    def case(name, a, b, tol=5e-3, **kw):
        ta, tb = MX.quantize_mx(a), MX.quantize_mx(b)
        da, db = MX.dequantize_mx(ta).float(), MX.dequantize_mx(tb).float()
        ra, _ = MX.quantize_mx_reference(a)
        report(f"mx_quant_{name}", da, ra, 1e-3)                     # kernel quantiser == restated quantiser
        got = MX.gemm_mx(ta, tb, **kw)
        torch.cuda.synchronize()
        ref = da @ db.transpose(-1, -2)
        return got, ref

    M, Nn, K = 256, 384, 256
    a = torch.randn(M, K, device=DEV, dtype=HALF)
    b = torch.randn(Nn, K, device=DEV, dtype=HALF)
    ones_a, ones_b = torch.ones_like(a), torch.ones_like(b)
    got, ref = case("uniform", ones_a, ones_b)
    report("mx_gemm_uniform", got, ref, 5e-3)
    rows = torch.pow(2.0, (torch.arange(M, device=DEV) % 11).float())[:, None]
    got, ref = case("a_rows", (ones_a.float() * rows).to(HALF), ones_b)
    report("mx_gemm_a_rows", got, ref, 5e-3)
    cols = torch.pow(2.0, (torch.arange(Nn, device=DEV) % 13).float())[:, None]
    got, ref = case("b_rows", ones_a, (ones_b.float() * cols).to(HALF))
    report("mx_gemm_b_rows", got, ref, 5e-3)
    kg = torch.pow(2.0, ((torch.arange(K, device=DEV) // 32) % 5).float())[None, :]
    got, ref = case("k_groups", (ones_a.float() * kg).to(HALF), (ones_b.float() / kg).to(HALF))
    report("mx_gemm_k_groups", got, ref, 5e-3)
    for (M, Nn, K) in [(128, 192, 128), (256, 384, 256), (384, 8192, 2048), (200, 328, 384), (4096, 2048, 8192)]:
        a = torch.randn(M, K, device=DEV, dtype=HALF) * torch.pow(
            2.0, torch.randint(-6, 6, (M, 1), device=DEV).float()).to(HALF)
        b = (torch.randn(Nn, K, device=DEV) * 0.05).to(HALF)
        got, ref = case(f"{M}x{Nn}x{K}", a, b)
        report(f"mx_gemm_{M}x{Nn}x{K}", got, ref, 5e-3)
        report_rms(f"mx_gemm_vs_bf16_{M}x{Nn}x{K}", got, a.float() @ b.float().t(), 6e-2)
    # epilogue: bias + gelu with the pre-activation copy, batched operands
    E_, M, Nn, K = 2, 256, 320, 256
    a = torch.randn(E_, M, K, device=DEV, dtype=HALF)
    b = (torch.randn(E_, Nn, K, device=DEV) * 0.05).to(HALF)
    got, ref = case("batched", a, b)
    report("mx_gemm_batched", got.reshape(E_ * M, Nn), ref.reshape(E_ * M, Nn), 5e-3)
    bias = torch.randn(Nn, device=DEV, dtype=HALF)
    aux = torch.empty(M, Nn, device=DEV, dtype=HALF)
    got, ref = case("gelu", a[0], b[0], bias=bias, gelu=True, aux=aux)
    report("mx_gemm_gelu_aux", aux, ref + bias.float(), 5e-3)
    report("mx_gemm_gelu", got, torch.nn.functional.gelu(ref + bias.float(), approximate="tanh"), 1e-2)
    a = torch.randn(4096, 2048, device=DEV, dtype=HALF)
    b = torch.randn(8192, 2048, device=DEV, dtype=HALF)
    ta, tb = MX.quantize_mx(a), MX.quantize_mx(b)
    ms = timeit(lambda: MX.gemm_mx(ta, tb))
    print(json.dumps({"name": "mx_gemm_perf_4096x8192x2048", "ms": ms,
                      "tflops": 2 * 4096 * 8192 * 2048 / ms / 1e9}), flush=True)
    ta2, tb2 = MX.quantize_mx(torch.randn(4096, 8192, device=DEV, dtype=HALF)), MX.quantize_mx(
        torch.randn(2048, 8192, device=DEV, dtype=HALF))
    ms = timeit(lambda: MX.gemm_mx(ta2, tb2))
    print(json.dumps({"name": "mx_gemm_perf_4096x2048x8192", "ms": ms,
                      "tflops": 2 * 4096 * 8192 * 2048 / ms / 1e9}), flush=True)
    ms = timeit(lambda: MX.quantize_mx(a))
    print(json.dumps({"name": "mx_quantize_4096x2048", "ms": ms, "gbs": 4096 * 2048 * 3 / ms / 1e6}), flush=True)
    # transposed quantiser: the same bytes and scales as transpose-copy + row quantiser, without the copy
    for shape in [(256, 384), (3, 384, 200), (8, 2048, 8192)]:
        w = (torch.randn(*shape, device=DEV) * 0.05).to(HALF)
        t1 = MX.quantize_mx_t(w)
        t0 = MX.quantize_mx(w.transpose(-1, -2).contiguous())
        same = bool(torch.equal(t1.q, t0.q)) and bool(torch.equal(t1.sf, t0.sf))
        rec = {"name": f"mx_quantize_t_{'x'.join(map(str, shape))}", "ok": same,
               "q_mismatch": int((t1.q != t0.q).sum()), "sf_mismatch": int((t1.sf != t0.sf).sum())}
        RESULTS.append(rec)
        print(json.dumps(rec), flush=True)
    w = (torch.randn(8, 2048, 8192, device=DEV) * 0.05).to(HALF)
    ms_t = timeit(lambda: MX.quantize_mx_t(w))
    ms_c = timeit(lambda: MX.quantize_mx(w.transpose(1, 2).contiguous()))
    print(json.dumps({"name": "mx_quantize_t_perf_8x2048x8192", "ms": ms_t, "ms_transpose_copy_then_quantize": ms_c,
                      "gbs": w.numel() * 3 / ms_t / 1e6}), flush=True)


GROUPS = {
    "mx": group_mx,
    "k5": group_k5,
    "attn_drop": group_attn_drop,
    "gemm2": group_gemm2,
    "fp8": group_fp8,
    "moe": group_moe,
    "detection": group_detection,
    "maskrcnn_step": group_maskrcnn_step,
    "attn_fwd": lambda: group_attn("fwd"),
    "attn_bwd": lambda: group_attn("bwd"),
    "attn_perf": group_attn_perf,
    "gemm_nt": lambda: group_gemm("nt"),
    "gemm_nn": lambda: group_gemm("nn"),
    "gemm_tn": lambda: group_gemm("tn"),
    "gemm_tt": lambda: group_gemm("tt"),
    "gemm_epi": group_gemm_epi,
    "gemm_perf": group_gemm_perf,
    "vocab_ce": group_vocab_ce,
    "norm": group_norm,
    "norm_perf": group_norm_perf,
    "optim": group_optim,
    "elementwise": group_elementwise,
}

try:  # optional groups registered by later modules
    from tools import kernel_check_ext  # type: ignore  # noqa: F401
except Exception:  # pragma: no cover
    pass


if __name__ == "__main__":
    names = sys.argv[1:] or list(GROUPS)
    t0 = time.time()
    for n in names:
        try:
            GROUPS[n]()
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            rec = {"name": n, "ok": False, "exception": repr(e)[:500]}
            RESULTS.append(rec)
            print(json.dumps(rec), flush=True)
    nbad = sum(1 for r in RESULTS if not r.get("ok"))
    print(json.dumps({"summary": names, "checks": len(RESULTS), "failed": nbad,
                      "seconds": round(time.time() - t0, 1)}), flush=True)
    sys.exit(1 if nbad else 0)
