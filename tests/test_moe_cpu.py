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


def _ep_worker(rank, world, port, q, top_k=1):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(5)
    full = MoELayer(32, 64, 4, dtype=torch.float32, top_k=top_k)        # all 4 experts, identical on both ranks
    ep = EPContext(dist.group.WORLD, world, rank)
    part = MoELayer(32, 64, 4, ep=ep, dtype=torch.float32, fused=False, top_k=top_k)
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


@pytest.mark.parametrize("top_k", [1, 2])
def test_expert_parallel_all_to_all_matches_local(top_k):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ep_worker, args=(r, 2, 29643 + top_k, q, top_k)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res


def _top2_reference(logits, C):
    """DeepSpeed `top2gating` (no noise) written the dense way: one-hot masks, cumsum locations,
    second choices placed after every first choice of the expert, capacity cut, renormalised gates."""
    T, E = logits.shape
    gates = logits.softmax(-1)
    i1 = gates.argmax(-1)
    m1 = torch.nn.functional.one_hot(i1, E)
    i2 = logits.masked_fill(m1.bool(), float("-inf")).argmax(-1)
    m2 = torch.nn.functional.one_hot(i2, E)
    loc1 = m1.cumsum(0) - 1
    loc2 = m2.cumsum(0) - 1 + m1.sum(0, keepdim=True)
    aux = (gates.mean(0) * m1.float().mean(0)).sum() * E
    m1, m2 = m1 * (loc1 < C), m2 * (loc2 < C)
    s1, s2 = (loc1 * m1).sum(1), (loc2 * m2).sum(1)
    g1, g2 = (gates * m1).sum(1), (gates * m2).sum(1)
    den = (g1 + g2).clamp(min=torch.finfo(torch.float32).eps)
    return (i1, i2), (m1.sum(1) > 0, m2.sum(1) > 0), (s1, s2), (g1 / den, g2 / den), aux


def test_top2_routing_matches_deepspeed_formulation():
    from b200flow.parallel.moe import gate_top2

    torch.manual_seed(3)
    T, E, C = 96, 4, 40            # 2T/E = 48 > C: some second choices are cut
    logits = torch.randn(T, E, requires_grad=True)
    gates, aux, expert, slot = gate_top2(logits, C)
    (i1, i2), (k1, k2), (s1, s2), (g1, g2), aux_ref = _top2_reference(logits.detach(), C)
    assert not bool(k2.all()) and bool(k1.all())
    for k, (i, keep, s, g) in enumerate(((i1, k1, s1, g1), (i2, k2, s2, g2))):
        assert torch.equal(expert[k][keep].long(), i[keep]) and (expert[k][~keep] == -1).all()
        assert torch.equal(slot[k][keep].long(), s[keep])
        assert torch.allclose(gates[k], g, atol=1e-6)
    assert abs(float(aux) - float(aux_ref)) < 1e-6
    # no (expert, slot) pair is handed out twice
    rows = (expert.long() * C + slot.long())[expert >= 0]
    assert rows.unique().numel() == rows.numel()
    (gates.sum() + aux).backward()
    assert logits.grad is not None and torch.isfinite(logits.grad).all()


def test_top2_layer_matches_dense_oracle_and_trains():
    from b200flow.parallel.moe import gate_top2

    torch.manual_seed(4)
    layer = MoELayer(32, 64, 4, dtype=torch.float32, capacity_factor=0.75, top_k=2)
    x = torch.randn(40, 32, requires_grad=True)
    y = layer(x)
    T = x.shape[0]
    with torch.no_grad():
        logits = torch.nn.functional.linear(x, layer.wg)[:, :layer.E].float()
        gates, _, expert, _ = gate_top2(logits, layer.capacity(T))
        ref = torch.zeros_like(x)
        for t in range(T):
            for k in range(2):
                e = int(expert[k, t])
                if e >= 0:
                    h = torch.nn.functional.gelu(torch.nn.functional.linear(x[t], layer.w1[e], layer.b1[e]),
                                                 approximate="tanh")
                    ref[t] += gates[k, t] * torch.nn.functional.linear(h, layer.w2[e], layer.b2[e])
    assert torch.allclose(y, ref, atol=1e-5)
    (y.square().sum() + layer.last_aux).backward()
    assert x.grad is not None and layer.wg.grad.abs().sum() > 0 and layer.w2.grad.abs().sum() > 0


def test_stock_library_twin_of_gpt_moe_trains_on_cpu():
    """`bench.py --impl twin --workload gpt-moe-1.3b` at toy size: the stock-ops MoE GPT runs forward / backward
    and its routing follows the same top-1 + capacity rule as the native router."""
    from b200flow.workloads.twin import TwinGPTMoE, _TwinMoE

    torch.manual_seed(0)
    m = TwinGPTMoE(256, 64, 4, 4, 128, 32, experts=4, moe_every=2, world=1, group=None, dtype=torch.float32,
                   device=torch.device("cpu"))
    t = torch.randint(0, 256, (2, 33))
    opt = torch.optim.AdamW(m.parameters(), lr=3e-3)
    losses = []
    for _ in range(8):
        loss = m(t[:, :-1], t[:, 1:])
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0] and len(m.expert_parameters()) == 2 * 4
    moe = _TwinMoE(32, 64, 4, 1, None, torch.float32, torch.device("cpu"))
    x = torch.randn(40, 32)
    y = moe(x)
    r = route_top1(moe.gate(x).float(), (max(4, 10) + 7) // 8 * 8)
    dropped = r["expert"] < 0
    assert torch.equal(y[dropped], torch.zeros_like(y[dropped])) and (y[~dropped].abs().sum(1) > 0).all()


def _ep_engine_worker(rank, world, port, q):
    import itertools

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    ps = None
    if world > 1:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from b200flow.parallel.groups import ParallelState

        ps = ParallelState.initialize(tp=1, pp=1, ep=world)
    torch.set_num_threads(1)
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import GPTModel, config_from_preset

    cfg = config_from_preset("gpt-moe-tiny")
    gas = 4 // world
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": gas,
                           "optimizer": {"params": {"lr": 2e-3, "weight_decay": 0.0}}}, ps, device=torch.device("cpu"), seed=4)
    torch.manual_seed(123)
    full = GPTModel(cfg, dtype=torch.float32, device=torch.device("cpu")).state_dict()
    with torch.no_grad():
        for name, p in eng.model.named_parameters():
            src = full[name]
            if getattr(p, "is_expert", False) and world > 1:
                el = p.shape[0]
                src = src[rank * el:(rank + 1) * el]
            p.copy_(src)
    for o in (eng.optimizer, eng.expert_optimizer):
        o.master.copy_(o.flat.params[o.lo:o.hi].float())
    g = torch.Generator().manual_seed(9)
    toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(4)]
    mine = toks[rank::world] if world > 1 else toks
    it = itertools.cycle([(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in mine])
    losses = []
    for _ in range(3):
        loss = eng.train_batch(it).clone()
        if world > 1:
            dist.all_reduce(loss)
            loss /= world
        losses.append(float(loss))
    if rank == 0:
        q.put(losses)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def test_engine_with_expert_parallelism_follows_the_single_process_curve():
    """GPT-MoE through the engine at EP = DP = 2 (experts split over the ranks, dense parameters under ZeRO-1 over
    both, expert parameters in their own group) against one process holding all experts on the same global batch."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    _ep_engine_worker(0, 1, 29661, q)
    want = q.get()
    procs = [ctx.Process(target=_ep_engine_worker, args=(r, 2, 29662, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got[-1] < got[0]
    for a, b in zip(got, want):
        assert abs(a - b) < 2e-3, (got, want)


def test_eval_capacity_factor_applies_in_eval_mode_only():
    layer = MoELayer(32, 64, 4, dtype=torch.float32, capacity_factor=1.0, eval_capacity_factor=2.0)
    assert layer.capacity(64) == 16
    layer.eval()
    assert layer.capacity(64) == 32
    layer.train()
    assert layer.capacity(64) == 16
    x = torch.randn(64, 32)
    layer.eval()
    y = layer(x)
    assert torch.isfinite(y).all()


def test_moe_with_pipeline_stages_is_refused():
    from b200flow.engines.ds_engine import TrainEngine
    from b200flow.models.gpt import config_from_preset

    class _Topo:
        tp, pp, dp, ep = 1, 2, 1, 1

    class _PS:
        topo, pp_rank, tp_rank, dp_rank = _Topo(), 0, 0, 0

    with pytest.raises(ValueError, match="pipeline"):
        TrainEngine(config_from_preset("gpt-moe-tiny"), {}, _PS(), device=torch.device("cpu"))
