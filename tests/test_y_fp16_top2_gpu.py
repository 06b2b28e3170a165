"""GPU checks of the fp16 loss-scale path and top-2 MoE routing (CPU twins: test_fp16_and_eval_cpu.py,
test_moe_cpu.py).  Both ride on kernels that have their own numerics tests — the sum-of-squares
kernel's overflow flag, the top-1 router, the dispatch / combine row movers — so what is checked here
is the composition on a device."""
import itertools

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


def test_top2_router_matches_cpu_and_layer_matches_fp32_oracle():
    from b200flow.parallel.moe import MoELayer, gate_top2, route_top2

    torch.manual_seed(8)
    T, E, C = 3000, 8, 512                      # 2T/E = 750 > C: second choices get cut
    logits = torch.randn(T, E)
    rg, rc = route_top2(logits.cuda(), C), route_top2(logits, C)
    assert torch.equal(rg["expert"].cpu(), rc["expert"]) and torch.equal(rg["slot"].cpu(), rc["slot"])
    assert int((rc["expert"][1] < 0).sum()) > 0

    layer = MoELayer(256, 512, 8, dtype=torch.bfloat16, device="cuda", top_k=2)
    x = (torch.randn(1024, 256, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_()
    y = layer(x)
    (y.float().square().sum() + layer.last_aux).backward()
    assert torch.isfinite(y).all() and torch.isfinite(x.grad).all() and torch.isfinite(layer.wg.grad).all()
    assert layer.w1.grad.abs().sum() > 0 and layer.wg.grad.abs().sum() > 0
    # fp32 oracle with the routing the device took (bf16 logits can flip near-ties against a CPU run)
    with torch.no_grad():
        xf = x.detach().float()
        lg = (xf @ layer.wg.float().t())[:, :8]
        gates, _, expert, _ = gate_top2(lg, layer.capacity(1024))
        ref = torch.zeros_like(xf)
        for k in range(2):
            for e in range(8):
                sel = (expert[k] == e).nonzero().flatten()
                if sel.numel():
                    h = torch.nn.functional.gelu(xf[sel] @ layer.w1[e].float().t() + layer.b1[e].float(), approximate="tanh")
                    ref[sel] += gates[k, sel, None] * (h @ layer.w2[e].float().t() + layer.b2[e].float())
    assert _rel(y, ref) < 6e-2


def test_fp16_loss_scaling_on_device_tracks_bf16_run_and_skips_overflow():
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    def engine(extra):
        cfg = config_from_preset("llama-tiny", hidden=256, heads=2, ffn=512, layers=2, max_seq=128)
        return cfg, initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
                                     "optimizer": {"params": {"lr": 1e-3}}, **extra}, device=torch.device("cuda"), seed=7)

    cfg, plain = engine({})
    _, scaled = engine({"fp16": {"enabled": True, "initial_scale_power": 12, "hysteresis": 1, "loss_scale_window": 2}})
    g = torch.Generator().manual_seed(0)
    data = [torch.randint(0, cfg.vocab_size, (2, 129), generator=g) for _ in range(2)]
    batches = [(t[:, :-1].contiguous().cuda(), t[:, 1:].contiguous().cuda()) for t in data]
    ia, ib = itertools.cycle(batches), itertools.cycle(batches)
    for _ in range(3):
        la, lb = float(plain.train_batch(ia)), float(scaled.train_batch(ib))
        assert abs(la - lb) < 2e-2, (la, lb)        # a power-of-two scale changes no rounding
    assert scaled.loss_scale == 8192.0 and scaled.skipped_steps == 0
    before, steps = scaled.flat.params.clone(), scaled.optimizer.step_count
    orig = scaled.flat.collect_autograd_grads

    def poisoned():
        orig()
        scaled.flat.grads[5] = float("nan")

    scaled.flat.collect_autograd_grads = poisoned
    scaled.train_batch(ib)
    scaled.flat.collect_autograd_grads = orig
    assert scaled.skipped_steps == 1 and scaled.loss_scale == 4096.0
    assert torch.equal(before, scaled.flat.params) and scaled.optimizer.step_count == steps
    after = float(scaled.train_batch(ib))                                        # still trains after the skip
    assert after == after and scaled.skipped_steps == 1 and scaled.optimizer.step_count == steps + 1
