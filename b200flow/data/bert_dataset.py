"""BERT pre-training samples (masked LM + next-sentence prediction) over an mmap corpus of sentences.

BASELINE.json config 2 is BERT-large pre-training; the reference's Megatron-DeepSpeed pin builds such samples
with its C++ `build_mapping` helper plus `build_training_sample` (SURVEY.md §2E "megatron/data/helpers.cpp").
The sentence-packing map is `b2f_build_bert_mapping` (csrc/host/data_helpers.cpp); what happens to one row of
that map is below:

* the sentences of a sample are split at a random point into segment A and segment B; with probability 1/2
  the segments are swapped and the sample is labelled "not next" (Megatron's `is_random` — segments come from
  the same document either way);
* the pair is truncated from the longer side, front or back at random, to fit `target - 3` tokens;
* `[CLS] A [SEP] B [SEP]`, token types 0 / 1, padding to max_seq_length;
* 15 % of the non-special positions are predicted: 80 % become `[MASK]`, 10 % a random token, 10 % stay.

The corpus format is the indexed dataset of `indexed_dataset.py` with one item per sentence and
`end_document()` after each document (what `tools/preprocess_data.py --split-sentences` writes).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .indexed_dataset import MMapIndexedDataset


def build_bert_mapping(docs: np.ndarray, sizes: np.ndarray, *, num_epochs: int, max_num_samples: int,
                       max_seq_length: int, short_seq_prob: float, seed: int, min_num_sent: int = 2,
                       long_sentence_len: int = 512) -> np.ndarray:
    """int64 [n, 3] rows of (first sentence, last sentence + 1, target length), shuffled."""
    from ..ops._lib import host_lib

    lib = host_lib()
    lib.b2f_build_bert_mapping.restype = ctypes.c_longlong
    docs = np.ascontiguousarray(docs, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)

    def call(buf):
        return lib.b2f_build_bert_mapping(
            docs.ctypes.data_as(ctypes.c_void_p), ctypes.c_longlong(len(docs) - 1),
            sizes.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(num_epochs), ctypes.c_longlong(max_num_samples),
            ctypes.c_int(max_seq_length), ctypes.c_double(short_seq_prob), ctypes.c_ulonglong(seed),
            ctypes.c_int(min_num_sent), ctypes.c_int(long_sentence_len),
            buf.ctypes.data_as(ctypes.c_void_p) if buf is not None else None)

    n = call(None)
    maps = np.empty((n, 3), dtype=np.int64)
    assert call(maps) == n
    return maps


class BertPretrainingDataset(torch.utils.data.Dataset):
    def __init__(self, prefix: str, *, vocab_size: int, cls_id: int, sep_id: int, mask_id: int, pad_id: int = 0,
                 max_seq_length: int = 512, num_samples: int | None = None, num_epochs: int | None = None,
                 masked_lm_prob: float = 0.15, short_seq_prob: float = 0.1, seed: int = 1234):
        self.ds = MMapIndexedDataset(prefix)
        assert len(self.ds.doc_idx) >= 2, "corpus has no document boundaries"
        self.vocab_size, self.cls, self.sep, self.mask, self.pad = vocab_size, cls_id, sep_id, mask_id, pad_id
        self.max_len, self.mlm_prob, self.seed = max_seq_length, masked_lm_prob, seed
        if num_epochs is None:
            num_epochs = 1 if num_samples is None else np.iinfo(np.int32).max - 1
        if num_samples is None:
            num_samples = np.iinfo(np.int64).max - 1
        self.maps = build_bert_mapping(np.asarray(self.ds.doc_idx), np.asarray(self.ds.sizes), num_epochs=num_epochs,
                                       max_num_samples=num_samples, max_seq_length=max_seq_length - 3,
                                       short_seq_prob=short_seq_prob, seed=seed)

    def __len__(self) -> int:
        return len(self.maps)

    def __getitem__(self, i: int) -> dict:
        first, last, target = (int(v) for v in self.maps[i])
        rng = np.random.RandomState((self.seed + i) % (2 ** 32))
        sents = [np.asarray(self.ds[s], dtype=np.int64) for s in range(first, last)]
        # segment split
        cut = 1 if len(sents) < 3 else int(rng.randint(1, len(sents)))
        a = np.concatenate(sents[:cut])
        b = np.concatenate(sents[cut:]) if len(sents) > 1 else np.zeros(0, dtype=np.int64)
        is_random = bool(rng.random_sample() < 0.5) and len(b) > 0
        if is_random:
            a, b = b, a
        # truncate the pair to the target length, trimming the longer side front or back at random
        while len(a) + len(b) > target:
            if len(a) > len(b):
                a = a[1:] if rng.random_sample() < 0.5 else a[:-1]
            else:
                b = b[1:] if rng.random_sample() < 0.5 else b[:-1]
        tokens = np.concatenate([[self.cls], a, [self.sep], b, [self.sep]])
        types = np.concatenate([np.zeros(len(a) + 2, dtype=np.int64), np.ones(len(b) + 1, dtype=np.int64)])
        # masked-LM positions
        cand = np.nonzero((tokens != self.cls) & (tokens != self.sep))[0]
        n_pred = max(1, int(round(len(tokens) * self.mlm_prob))) if len(cand) else 0
        picked = np.sort(rng.permutation(cand)[:n_pred])
        labels = np.full(self.max_len, -100, dtype=np.int64)
        labels[picked] = tokens[picked]
        roll = rng.random_sample(len(picked))
        masked = tokens.copy()
        masked[picked[roll < 0.8]] = self.mask
        rnd = picked[(roll >= 0.8) & (roll < 0.9)]
        masked[rnd] = rng.randint(0, self.vocab_size, size=len(rnd))
        n = len(tokens)
        out_tokens = np.full(self.max_len, self.pad, dtype=np.int64)
        out_types = np.zeros(self.max_len, dtype=np.int64)
        attn = np.zeros(self.max_len, dtype=np.int64)
        out_tokens[:n], out_types[:n], attn[:n] = masked, types, 1
        return {"input_ids": torch.from_numpy(out_tokens), "token_type_ids": torch.from_numpy(out_types),
                "attention_mask": torch.from_numpy(attn), "labels": torch.from_numpy(labels),
                "next_sentence_label": torch.tensor(int(is_random))}
