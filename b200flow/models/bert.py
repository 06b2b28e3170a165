"""BERT (base / large) on the b200flow ops: sequence classification and pre-training heads.

The reference fine-tunes `bert-base-cased` through Accelerate's `complete_nlp_example.py`
(examples/accelerate/bert-glue-mrpc/pretrain.yaml:43-51) and through Ray Train + Lightning
(examples/ray/lightning-bert/fine-tune.yaml:46-49); BASELINE.json config 2 pre-trains
BERT-large.  Architecture is the original post-LN encoder; every matmul is the tcgen05 GEMM
(bias / GELU in its epilogue), attention is the tcgen05 flash kernel with key-padding
lengths, LayerNorm fuses the residual add.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
from torch import nn

from ..ops import attention as A
from ..ops import elementwise as E
from ..ops import gemm as G
from ..ops import norm as N


@dataclass
class BertConfig:
    vocab_size: int = 28996          # bert-base-cased
    hidden: int = 768
    layers: int = 12
    heads: int = 12
    ffn: int = 3072
    max_pos: int = 512
    type_vocab: int = 2
    dropout: float = 0.1
    eps: float = 1e-12
    num_labels: int = 2
    init_std: float = 0.02

    def param_count(self) -> int:
        per = 4 * self.hidden * self.hidden + 2 * self.hidden * self.ffn + 9 * self.hidden + self.ffn
        emb = (self.vocab_size + self.max_pos + self.type_vocab + 2) * self.hidden
        return self.layers * per + emb


PRESETS = {
    "bert-base": dict(),
    "bert-large": dict(vocab_size=30592,   # 30522 padded to a multiple of 128 (Megatron-style)
                        hidden=1024, layers=24, heads=16, ffn=4096),
    "bert-tiny": dict(vocab_size=1024, hidden=128, layers=2, heads=2, ffn=256, max_pos=128),
}


def bert_config(name: str, **over) -> BertConfig:
    kw = dict(PRESETS[name])
    kw.update(over)
    return BertConfig(**kw)


def _p(*shape, std=None, ones=False, dtype=None, device=None):
    t = torch.empty(*shape, dtype=dtype, device=device)
    if ones:
        t.fill_(1.0)
    elif std is None:
        t.zero_()
    else:
        t.normal_(0, std)
    return nn.Parameter(t)


class BertLayer(nn.Module):
    def __init__(self, cfg: BertConfig, dtype, device):
        super().__init__()
        self.cfg = cfg
        H, F, kw = cfg.hidden, cfg.ffn, dict(dtype=dtype, device=device)
        s = cfg.init_std
        self.wqkv, self.bqkv = _p(3 * H, H, std=s, **kw), _p(3 * H, **kw)
        self.wo, self.bo = _p(H, H, std=s, **kw), _p(H, **kw)
        self.ln1_w, self.ln1_b = _p(H, ones=True, **kw), _p(H, **kw)
        self.w1, self.b1 = _p(F, H, std=s, **kw), _p(F, **kw)
        self.w2, self.b2 = _p(H, F, std=s, **kw), _p(H, **kw)
        self.ln2_w, self.ln2_b = _p(H, ones=True, **kw), _p(H, **kw)

    def forward(self, x, kv_lens, seed):
        cfg = self.cfg
        b, s, H = x.shape
        nh, d = cfg.heads, H // cfg.heads
        p = cfg.dropout if self.training else 0.0
        qkv = G.linear(x, self.wqkv, self.bqkv).view(b, s, 3, nh, d)
        o = A.flash_attention_qkvpacked(qkv, causal=False, scale=1.0 / math.sqrt(d), kv_lens=kv_lens)
        a = G.linear(o.reshape(b, s, H), self.wo, self.bo)
        if p > 0:
            a = E.bias_dropout_add(a, None, None, p, seed)
        x = _post_ln(a, x, self.ln1_w, self.ln1_b, cfg.eps)
        f = G.mlp_gelu(x, self.w1, self.b1, self.w2, self.b2)
        if p > 0:
            f = E.bias_dropout_add(f, None, None, p, seed + 1)
        return _post_ln(f, x, self.ln2_w, self.ln2_b, cfg.eps)


def _post_ln(sub, residual, w, b, eps):
    out = N.norm(sub, w, b, residual=residual, eps=eps, rms=False)
    return out[0]  # (normed, pre-norm sum)


class BertModel(nn.Module):
    def __init__(self, cfg: BertConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.cfg = cfg
        kw = dict(dtype=dtype, device=device)
        H = cfg.hidden
        self.word = _p(cfg.vocab_size, H, std=cfg.init_std, **kw)
        self.pos = _p(cfg.max_pos, H, std=cfg.init_std, **kw)
        self.tok_type = _p(cfg.type_vocab, H, std=cfg.init_std, **kw)
        self.emb_ln_w, self.emb_ln_b = _p(H, ones=True, **kw), _p(H, **kw)
        self.layers = nn.ModuleList([BertLayer(cfg, dtype, device) for _ in range(cfg.layers)])
        self._step = 0

    def forward(self, input_ids, token_type_ids=None, attention_mask=None):
        b, s = input_ids.shape
        emb = torch.nn.functional.embedding(input_ids, self.word) + self.pos[:s].unsqueeze(0)
        if token_type_ids is not None:
            emb = emb + torch.nn.functional.embedding(token_type_ids, self.tok_type)
        else:
            emb = emb + self.tok_type[0]
        x = N.norm(emb.contiguous(), self.emb_ln_w, self.emb_ln_b, eps=self.cfg.eps, rms=False)
        # right-padded batches: attention_mask -> valid key length per sample
        kv_lens = attention_mask.sum(-1).to(torch.int32) if attention_mask is not None else None
        self._step += 1
        for i, layer in enumerate(self.layers):
            x = layer(x, kv_lens, self._step * 1000 + i * 2)
        return x


class BertForSequenceClassification(nn.Module):
    def __init__(self, cfg: BertConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.cfg = cfg
        self.bert = BertModel(cfg, dtype, device)
        kw = dict(dtype=dtype, device=device)
        self.pool_w, self.pool_b = _p(cfg.hidden, cfg.hidden, std=cfg.init_std, **kw), _p(cfg.hidden, **kw)
        self.cls_w, self.cls_b = _p(cfg.num_labels, cfg.hidden, std=cfg.init_std, **kw), _p(cfg.num_labels, **kw)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None):
        x = self.bert(input_ids, token_type_ids, attention_mask)
        pooled = torch.tanh(torch.nn.functional.linear(x[:, 0], self.pool_w, self.pool_b))
        logits = torch.nn.functional.linear(pooled, self.cls_w, self.cls_b).float()
        loss = torch.nn.functional.cross_entropy(logits, labels) if labels is not None else None
        return {"loss": loss, "logits": logits}


class BertForPreTraining(nn.Module):
    """Masked-LM + next-sentence heads (BASELINE.json config 2)."""

    def __init__(self, cfg: BertConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.cfg = cfg
        self.bert = BertModel(cfg, dtype, device)
        kw = dict(dtype=dtype, device=device)
        H = cfg.hidden
        self.tr_w, self.tr_b = _p(H, H, std=cfg.init_std, **kw), _p(H, **kw)
        self.tr_ln_w, self.tr_ln_b = _p(H, ones=True, **kw), _p(H, **kw)
        self.dec_b = _p(cfg.vocab_size, **kw)
        self.bert.word.ddp_uses = 2     # embedding lookup + tied decoder
        self.pool_w, self.pool_b = _p(H, H, std=cfg.init_std, **kw), _p(H, **kw)
        self.nsp_w, self.nsp_b = _p(2, H, std=cfg.init_std, **kw), _p(2, **kw)
        self.gather_masked = True

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None,
                next_sentence_label=None, masked_lm_positions=None):
        """`masked_lm_positions` (optional, int64 [n_masked]): flat indices into [batch*seq] of the
        positions whose label is not -100, prepared by the collator.  The MLM transform, the tied
        decoder GEMM and the loss then run on those rows only (same loss; the other rows have zero
        gradient anyway).  Without it the positions are found on the device (one host sync)."""
        cfg = self.cfg
        x = self.bert(input_ids, token_type_ids, attention_mask)
        b, s, H = x.shape
        flat_labels = labels.reshape(-1)
        if masked_lm_positions is None and x.is_cuda and self.gather_masked:
            masked_lm_positions = (flat_labels != -100).nonzero().squeeze(-1)
        xm = x.reshape(b * s, H)
        if masked_lm_positions is not None:
            n_m = masked_lm_positions.numel()
            pad = (-n_m) % 8          # keep GEMM rows a multiple of 8; padded rows carry label -100
            if pad:
                masked_lm_positions = torch.cat([masked_lm_positions, masked_lm_positions.new_zeros(pad)])
            xm = xm.index_select(0, masked_lm_positions)
            flat_labels = flat_labels.index_select(0, masked_lm_positions)
            if pad:
                flat_labels = flat_labels.clone()
                flat_labels[n_m:] = -100
        h = G.linear_gelu(xm, self.tr_w, self.tr_b)
        h = N.norm(h, self.tr_ln_w, self.tr_ln_b, eps=cfg.eps, rms=False)
        logits = G.linear(h, self.bert.word, self.dec_b)   # tied decoder
        mlm = E.cross_entropy(logits, flat_labels, ignore_index=-100)
        pooled = torch.tanh(torch.nn.functional.linear(x[:, 0], self.pool_w, self.pool_b))
        nsp_logits = torch.nn.functional.linear(pooled, self.nsp_w, self.nsp_b).float()
        loss = mlm
        if next_sentence_label is not None:
            loss = loss + torch.nn.functional.cross_entropy(nsp_logits, next_sentence_label)
        return {"loss": loss, "mlm_loss": mlm.detach(), "nsp_logits": nsp_logits}
