"""DeepSpeed scheduler types of a ds_config (b200flow/parallel/lr_schedules.py)."""
import math
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.parallel.lr_schedules import make_schedule  # noqa: E402


def test_warmup_lr_log_and_linear():
    f = make_schedule({"type": "WarmupLR", "params": {"warmup_min_lr": 0.0, "warmup_max_lr": 1e-3, "warmup_num_steps": 100}}, 5e-4)
    assert f(0) == 0.0 and abs(f(50) - 1e-3 * math.log(51) / math.log(101)) < 1e-12 and f(100) == 1e-3 and f(10 ** 6) == 1e-3
    g = make_schedule({"type": "WarmupLR", "params": {"warmup_max_lr": 1e-3, "warmup_num_steps": 100, "warmup_type": "linear"}}, 5e-4)
    assert abs(g(25) - 2.5e-4) < 1e-15


def test_warmup_decay_and_cosine():
    f = make_schedule({"type": "WarmupDecayLR", "params": {"warmup_max_lr": 1e-3, "warmup_num_steps": 10, "total_num_steps": 110,
                                                           "warmup_type": "linear"}}, 1e-3)
    assert abs(f(5) - 5e-4) < 1e-15 and f(10) == 1e-3 and abs(f(60) - 5e-4) < 1e-15 and f(110) == 0.0 and f(500) == 0.0
    g = make_schedule({"type": "WarmupCosineLR", "params": {"total_num_steps": 110, "warmup_num_steps": 10, "cos_min_ratio": 0.1}}, 2e-3)
    assert g(0) == 0.0 and abs(g(5) - 1e-3) < 1e-15 and abs(g(10) - 2e-3) < 1e-15
    assert abs(g(60) - 2e-3 * (0.1 + 0.9 * 0.5)) < 1e-12 and abs(g(110) - 2e-4) < 1e-12 and abs(g(999) - 2e-4) < 1e-12
    with pytest.raises(ValueError):
        make_schedule({"type": "OneCycle", "params": {}}, 1e-3)
    assert make_schedule(None, 1e-3) is None and make_schedule({"params": {}}, 1e-3) is None


def test_megatron_block_keeps_working():
    f = make_schedule({"params": {"max_lr": 1e-3, "min_lr": 1e-5, "warmup_steps": 10, "decay_steps": 110, "decay_style": "linear"}}, 0.0)
    assert f(5) == 5e-4 and abs(f(60) - (1e-5 + 0.5 * (1e-3 - 1e-5))) < 1e-12


def test_engine_follows_a_deepspeed_scheduler_block():
    import itertools

    import torch

    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=2, vocab_size=128, max_seq=32)
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "optimizer": {"params": {"lr": 1e-3}},
                           "scheduler": {"type": "WarmupDecayLR", "params": {"warmup_max_lr": 1e-3, "warmup_num_steps": 2,
                                                                           "total_num_steps": 6, "warmup_type": "linear"}}},
                     device=torch.device("cpu"))
    t = torch.randint(0, 128, (2, 33))
    it = itertools.cycle([(t[:, :-1].contiguous(), t[:, 1:].contiguous())])
    seen = []
    for _ in range(5):
        eng.train_batch(it)
        seen.append(eng.optimizer.lr)
    assert seen == pytest.approx([5e-4, 1e-3, 7.5e-4, 5e-4, 2.5e-4])


def test_train_batch_size_rule_of_the_ds_config():
    import torch

    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=2, vocab_size=128, max_seq=32)
    mk = lambda ds: initialize(cfg, ds, device=torch.device("cpu"))  # noqa: E731
    assert mk({"train_batch_size": 8, "train_micro_batch_size_per_gpu": 2}).gas == 4           # accumulation derived
    assert mk({"train_batch_size": 6, "gradient_accumulation_steps": 3}).micro_batch == 2      # micro batch derived
    e = mk({"train_batch_size": 8, "train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 4})
    assert (e.micro_batch, e.gas) == (2, 4)
    with pytest.raises(ValueError, match="train_batch_size"):
        mk({"train_batch_size": 10, "train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 4})


def test_deepspeed_accessors_and_tuple_form():
    import torch

    from b200flow.engines.ds_engine import initialize_tuple
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=2, vocab_size=128, max_seq=32)
    eng, opt, loader, sched = initialize_tuple(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 3,
                                                      "optimizer": {"params": {"lr": 1e-3}}}, device=torch.device("cpu"))
    assert opt is eng.optimizer and loader is None and sched.get_lr() == [1e-3]
    assert (eng.train_micro_batch_size_per_gpu(), eng.gradient_accumulation_steps(), eng.train_batch_size()) == (2, 3, 6)
    assert eng.module is eng.model and eng.eval().model.training is False and eng.train().model.training is True
