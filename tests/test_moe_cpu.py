"""MoE routing math and expert parallelism on CPU/gloo against dense per-token oracles."""
import os
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.parallel.moe import EPContext, MoELayer, route_top1  # noqa: E402


def test_router_capacity_and_aux_terms():
    torch.manual_seed(0)
    T, E, C = 64, 4, 12
    logits = torch.randn(T, E)
    r = route_top1(logits, C)
    probs = logits.softmax(-1)
    best = probs.argmax(-1)
    for e in range(E):
        idx = (best == e).nonzero().flatten()
        kept = idx[:C]
        assert (r["expert"][kept] == e).all() and (r["slot"][kept] == torch.arange(len(kept))).all()
        assert (r["expert"][idx[C:]] == -1).all()            # over-capacity tokens are dropped
        assert int(r["counts"][e]) == min(C, len(idx))
        assert abs(float(r["ce"][e]) - len(idx)) < 1e-6
    assert torch.allclose(r["me"], probs.sum(0), atol=1e-5)
    assert torch.allclose(r["gate"][r["expert"] >= 0], probs.max(-1).values[r["expert"] >= 0])


def _dense_oracle(layer, x):
    """Per-token loop: y[t] = gate * FFN_{expert(t)}(x[t]) for kept tokens."""
    T = x.shape[0]
    logits = torch.nn.functional.linear(x, layer.wg)[:, :layer.E].float()
    r = route_top1(logits, layer.capacity(T))
    y = torch.zeros_like(x)
    for t in range(T):
        e = int(r["expert"][t])
        if e >= 0:
            h = torch.nn.functional.gelu(torch.nn.functional.linear(x[t], layer.w1[e], layer.b1[e]), approximate="tanh")
            y[t] = r["gate"][t] * torch.nn.functional.linear(h, layer.w2[e], layer.b2[e])
    return y


def test_single_process_layer_matches_dense_oracle():
    torch.manual_seed(1)
    layer = MoELayer(32, 64, 4, dtype=torch.float32, capacity_factor=1.0)
    x = torch.randn(40, 32, requires_grad=True)
    y = layer(x)
    ref = _dense_oracle(layer, x.detach())
    assert torch.allclose(y, ref, atol=1e-5)
    (y.sum() + layer.last_aux).backward()
    assert x.grad is not None and layer.wg.grad is not None and layer.w1.grad.abs().sum() > 0


def _ep_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(5)
    full = MoELayer(32, 64, 4, dtype=torch.float32)                     # all 4 experts, identical on both ranks
    ep = EPContext(dist.group.WORLD, world, rank)
    part = MoELayer(32, 64, 4, ep=ep, dtype=torch.float32, fused=False)
    with torch.no_grad():
        part.wg.copy_(full.wg)
        lo = rank * part.El
        for a, b in ((part.w1, full.w1), (part.b1, full.b1), (part.w2, full.w2), (part.b2, full.b2)):
            a.copy_(b[lo:lo + part.El])
    torch.manual_seed(100 + rank)
    x = torch.randn(48, 32)
    xa, xb = x.clone().requires_grad_(), x.clone().requires_grad_()
    ya = part(xa)
    yb = full(xb)
    ya.square().sum().backward()
    yb.square().sum().backward()
    ok = torch.allclose(ya, yb, atol=1e-5) and torch.allclose(xa.grad, xb.grad, atol=1e-4)
    # expert grads on the owner = sum over BOTH ranks' tokens of the full layer's grads for that expert
    g = full.w1.grad.clone()
    dist.all_reduce(g)
    ok = ok and torch.allclose(part.w1.grad, g[lo:lo + part.El], atol=1e-4)
    q.put((rank, bool(ok), float((ya - yb).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


def test_expert_parallel_all_to_all_matches_local():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ep_worker, args=(r, 2, 29641, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
