"""GPU: BERT on the native kernels (fused FFN, key-length attention, group LayerNorm) against the
fp32 CPU oracle, and the masked-position MLM head against the all-positions head."""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(cfg, b=4, s=128, seed=0):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(10, cfg.vocab_size - 10, (b, s), generator=g)
    labels = torch.full((b, s), -100, dtype=torch.long)
    m = torch.rand(b, s, generator=g) < 0.15
    labels[m] = ids[m]
    lens = torch.tensor([s, s - 17, s // 2, s - 1])[:b]
    mask = (torch.arange(s)[None, :] < lens[:, None]).long()
    labels[mask == 0] = -100
    return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": mask,
            "labels": labels, "next_sentence_label": torch.randint(0, 2, (b,), generator=g)}


def test_bert_pretraining_loss_and_grads_match_cpu_oracle():
    from b200flow.models.bert import BertForPreTraining, bert_config

    cfg = bert_config("bert-tiny", dropout=0.0)
    torch.manual_seed(3)
    cpu = BertForPreTraining(cfg, dtype=torch.float32, device="cpu")
    gpu = BertForPreTraining(cfg, dtype=torch.bfloat16, device="cuda")
    gpu.load_state_dict({k: v.to(torch.bfloat16) for k, v in cpu.state_dict().items()})
    cpu.load_state_dict({k: v.float() for k, v in gpu.state_dict().items()})   # same rounded weights
    batch = _batch(cfg)
    lc = cpu(**batch)["loss"]
    lg = gpu(**{k: v.cuda() for k, v in batch.items()})["loss"]
    assert abs(float(lc) - float(lg)) < 0.05, (float(lc), float(lg))
    lc.backward()
    lg.backward()
    for (n, pc), (_, pg) in zip(cpu.named_parameters(), gpu.named_parameters()):
        if pc.grad is None:
            continue
        ref = pc.grad.float()
        got = pg.grad.float().cpu()
        err = (got - ref).abs().max() / (ref.abs().max() + 1e-6)
        assert err < 0.08, (n, float(err))


def test_masked_position_head_equals_all_position_head():
    from b200flow.models.bert import BertForPreTraining, bert_config

    cfg = bert_config("bert-tiny", dropout=0.0)
    torch.manual_seed(5)
    model = BertForPreTraining(cfg, dtype=torch.bfloat16, device="cuda")
    batch = {k: v.cuda() for k, v in _batch(cfg, seed=4).items()}
    model.gather_masked = True
    l1 = model(**batch)["loss"]
    pos = (batch["labels"].reshape(-1) != -100).nonzero().squeeze(-1)
    l2 = model(**batch, masked_lm_positions=pos)["loss"]
    model.gather_masked = False
    l3 = model(**batch)["loss"]
    assert abs(float(l1) - float(l2)) < 1e-3
    assert abs(float(l1) - float(l3)) < 2e-2, (float(l1), float(l3))
