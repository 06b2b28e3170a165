"""Optimizer-state handling on CPU: the fused AdamW front end keeps fp32 master weights and moments across a
state-dict round trip with bf16 parameters, and the Accelerate facade trains with a stock torch optimizer."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200flow.optim import AdamW


def _run(dtype, steps, opt=None, model=None):
    torch.manual_seed(0)
    if model is None:
        model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1)).to(dtype)
        opt = AdamW(model.parameters(), lr=1e-2, weight_decay=0.01)
    g = torch.Generator().manual_seed(1)
    for _ in range(steps):
        x, y = torch.randn(16, 8, generator=g).to(dtype), torch.randn(16, 1, generator=g).to(dtype)
        torch.nn.functional.mse_loss(model(x), y).backward()
        opt.step()
        opt.zero_grad()
    return model, opt


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_adamw_state_stays_fp32_through_load_state_dict(dtype):
    model, opt = _run(dtype, 3)
    sd = opt.state_dict()
    weights = {k: v.clone() for k, v in model.state_dict().items()}
    torch.manual_seed(5)
    model2 = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1)).to(dtype)
    model2.load_state_dict(weights)
    opt2 = AdamW(model2.parameters(), lr=1e-2, weight_decay=0.01)
    opt2.load_state_dict(sd)
    for p, q in zip(model.parameters(), model2.parameters()):
        for k in ("master", "exp_avg", "exp_avg_sq"):
            assert opt2.state[q][k].dtype == torch.float32, (k, opt2.state[q][k].dtype)
            assert torch.equal(opt2.state[q][k], opt.state[p][k])          # bit-exact, not rounded through bf16
    assert opt2._step == 3
    # both continue identically (same data stream position is re-created by the fixed generator seed)
    g = torch.Generator().manual_seed(77)
    x, y = torch.randn(16, 8, generator=g).to(dtype), torch.randn(16, 1, generator=g).to(dtype)
    for m, o in ((model, opt), (model2, opt2)):
        torch.nn.functional.mse_loss(m(x), y).backward()
        o.step()
    for p, q in zip(model.parameters(), model2.parameters()):
        assert torch.equal(p, q)


def test_multi_tensor_table_rejects_non_fp32_state():
    from b200flow.ops.optim import MultiTensorTable

    class _NoLib:       # the table is built before any kernel runs; stub the one library query it makes
        pass

    import b200flow.ops._lib as _lib

    orig = _lib.call_int
    _lib.call_int = lambda name, *a: 65536
    try:
        ok = [torch.zeros(8)]
        bad = [torch.zeros(8, dtype=torch.bfloat16)]
        with pytest.raises(AssertionError):
            MultiTensorTable(bad, ok, ok, ok, None)
        with pytest.raises(AssertionError):
            MultiTensorTable(ok, ok, bad, ok, None)
        MultiTensorTable(ok, ok, ok, ok, None)
    finally:
        _lib.call_int = orig


def _stock_opt_worker(rank, world, port, kind, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from b200flow.engines.accelerate_api import Accelerator

    acc = Accelerator(cpu=True)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1))
    opt = torch.optim.SGD(model.parameters(), lr=0.05) if kind == "sgd" else torch.optim.AdamW(model.parameters(), lr=1e-2)
    model, opt = acc.prepare(model, opt)
    g = torch.Generator().manual_seed(1)
    x, y = torch.randn(3, 8, 8, generator=g), torch.randn(3, 8, 1, generator=g)
    for step in range(3):
        lo = rank * 4
        loss = torch.nn.functional.mse_loss(model(x[step, lo:lo + 4]), y[step, lo:lo + 4])
        acc.backward(loss)
        opt.step()
        opt.zero_grad()
    if rank == 0:
        q.put(torch.cat([p.detach().reshape(-1) for p in model.parameters()]).tolist())
    acc.wait_for_everyone()
    dist.destroy_process_group()


@pytest.mark.parametrize("kind,port", [("sgd", 29651), ("adamw", 29652)])
def test_stock_torch_optimizer_trains_through_the_accelerator(kind, port):
    """`accelerator.prepare(torch.optim.AdamW(...))` is what the reference's example does: the DDP hooks park the
    gradients in the flat fp32 buffer, and the wrapper must hand them to an optimizer that only reads `p.grad`."""
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1))
    start = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).clone()
    opt = torch.optim.SGD(model.parameters(), lr=0.05) if kind == "sgd" else torch.optim.AdamW(model.parameters(), lr=1e-2)
    g = torch.Generator().manual_seed(1)
    x, y = torch.randn(3, 8, 8, generator=g), torch.randn(3, 8, 1, generator=g)
    for step in range(3):
        torch.nn.functional.mse_loss(model(x[step]), y[step]).backward()
        opt.step()
        opt.zero_grad()
    want = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_stock_opt_worker, args=(r, 2, port, kind, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = torch.tensor(q.get(timeout=300))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert not torch.allclose(got, start)                       # it trained
    assert torch.allclose(got, want, atol=1e-5), (got - want).abs().max()
