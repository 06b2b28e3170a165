"""GPU end-to-end: the engine's loss curve on the native kernels tracks the fp32 CPU oracle."""
import itertools

import pytest
import torch

pytestmark = pytest.mark.gpu


def _curve(device, steps=5):
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=256, heads=2, ffn=512, layers=2, max_seq=128)
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
                           "optimizer": {"params": {"lr": 1e-3}}}, device=torch.device(device), seed=7)
    g = torch.Generator().manual_seed(0)
    data = [torch.randint(0, cfg.vocab_size, (2, 129), generator=g) for _ in range(2)]
    batches = [(t[:, :-1].contiguous().to(device), t[:, 1:].contiguous().to(device)) for t in data]
    return [float(eng.train_batch(itertools.cycle(batches))) for _ in range(steps)], eng


def test_loss_parity_with_cpu_oracle():
    gpu, eng = _curve("cuda")
    cpu, _ = _curve("cpu")
    assert gpu[-1] < gpu[0]
    for a, b in zip(gpu, cpu):
        assert abs(a - b) < 0.15, (gpu, cpu)


def test_gpt2_family_runs():
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("gpt2-tiny")
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "optimizer": {"params": {"lr": 1e-3}}})
    t = torch.randint(0, cfg.vocab_size, (2, 129), device="cuda")
    it = itertools.cycle([(t[:, :-1].contiguous(), t[:, 1:].contiguous())])
    l0 = float(eng.train_batch(it))
    for _ in range(5):
        l1 = float(eng.train_batch(it))
    assert l1 < l0
