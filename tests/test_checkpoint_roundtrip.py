"""Megatron-DeepSpeed checkpoint layout round trips on CPU: file names, `latest`, and a
restored engine that continues exactly like the one that saved — dense GPT and GPT-MoE (expert files)."""
import itertools

import pytest
import torch

from b200flow.engines.ds_engine import initialize
from b200flow.models.gpt import config_from_preset


def _engine(preset, seed=11):
    cfg = config_from_preset(preset)
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1,
                           "optimizer": {"params": {"lr": 1e-3}}}, device=torch.device("cpu"), seed=seed)
    g = torch.Generator().manual_seed(5)
    toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(3)]
    batches = [(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks]
    return cfg, eng, batches


@pytest.mark.parametrize("preset,expert_files", [("gpt2-tiny", False), ("gpt-moe-tiny", True)])
def test_save_load_continues_identically(tmp_path, preset, expert_files):
    cfg, eng, batches = _engine(preset)
    it = itertools.cycle(batches)
    for _ in range(2):
        eng.train_batch(it)
    eng.save_checkpoint(str(tmp_path))
    assert (tmp_path / "latest").read_text() == "global_step2"
    assert (tmp_path / "latest_checkpointed_iteration.txt").read_text().strip() == "2"
    names = {p.name for p in (tmp_path / "global_step2").iterdir()}
    assert "mp_rank_00_model_states.pt" in names
    assert any(n.endswith("zero_pp_rank_0_mp_rank_00_optim_states.pt") for n in names)
    if expert_files:
        assert any(n.startswith("layer_") and "_expert_0_mp_rank_00_model_states.pt" in n for n in names), names
        assert "expp_rank_0_mp_rank_00_optim_states.pt" in names
    after_save = [float(eng.train_batch(it)) for _ in range(2)]

    _, fresh, _ = _engine(preset, seed=99)          # different init: everything must come from the files
    path, client = fresh.load_checkpoint(str(tmp_path))
    assert fresh.global_steps == 2 and path.endswith("global_step2")
    it2 = itertools.cycle(batches)
    for _ in range(2):
        next(it2)                                    # the data position is the caller's business
    resumed = [float(fresh.train_batch(it2)) for _ in range(2)]
    assert resumed == pytest.approx(after_save, rel=1e-5, abs=1e-6)
