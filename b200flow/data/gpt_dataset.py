"""GPT pre-training dataset over an mmap indexed corpus: document order, sample cut points and
sample shuffle — the three index arrays Megatron's `GPTDataset` builds with its C++ helper
(SURVEY.md §2H C31, §3.1 "build_train_valid_test_data_iterators").  The cut-point walk runs in
`csrc/host/data_helpers.cpp` (`b2f_build_sample_idx`); arrays are cached next to the data as
`.npy` like Megatron's `--data-cache-path`.
"""
from __future__ import annotations

import ctypes
import hashlib
from pathlib import Path

import numpy as np
import torch

from .indexed_dataset import MMapIndexedDataset


def parse_split(split: str, n: int) -> list[tuple[int, int]]:
    """`--split 949,50,1` -> document ranges for train / valid / test."""
    w = [float(x) for x in split.replace("/", ",").split(",")]
    w += [0.0] * (3 - len(w))
    tot = sum(w)
    bounds, acc = [0], 0.0
    for x in w[:3]:
        acc += x / tot
        bounds.append(int(round(acc * n)))
    bounds[-1] = n
    return [(bounds[i], bounds[i + 1]) for i in range(3)]


def _host():
    from ..ops._lib import host_lib

    lib = host_lib()
    lib.b2f_build_sample_idx.restype = ctypes.c_longlong
    return lib


def build_indices(sizes: np.ndarray, documents: np.ndarray, seq_length: int, num_samples: int,
                  seed: int):
    lib = _host()
    tokens_per_epoch = int(sizes[documents].astype(np.int64).sum())
    assert tokens_per_epoch > seq_length, "corpus smaller than one sample"
    epochs = 1
    while (epochs * tokens_per_epoch - 1) // seq_length < num_samples:
        epochs += 1
    doc_idx = np.empty(len(documents) * epochs, dtype=np.int32)
    docs32 = np.ascontiguousarray(documents, dtype=np.int32)
    lib.b2f_build_doc_idx(doc_idx.ctypes.data_as(ctypes.c_void_p), docs32.ctypes.data_as(ctypes.c_void_p),
                          ctypes.c_longlong(len(docs32)), ctypes.c_int(epochs), ctypes.c_ulonglong(seed))
    total = (epochs * tokens_per_epoch - 1) // seq_length
    sample_idx = np.zeros((total + 1, 2), dtype=np.int64)
    sizes32 = np.ascontiguousarray(sizes, dtype=np.int32)
    got = lib.b2f_build_sample_idx(sizes32.ctypes.data_as(ctypes.c_void_p),
                                   doc_idx.ctypes.data_as(ctypes.c_void_p), ctypes.c_longlong(len(doc_idx)),
                                   ctypes.c_int(seq_length), ctypes.c_longlong(total),
                                   sample_idx.ctypes.data_as(ctypes.c_void_p))
    assert got == total + 1, f"sample index walk ended early ({got} of {total + 1})"
    shuffle_idx = np.empty(total, dtype=np.int64)
    lib.b2f_build_shuffle_idx(shuffle_idx.ctypes.data_as(ctypes.c_void_p), ctypes.c_longlong(total),
                              ctypes.c_ulonglong(seed + 1))
    return doc_idx, sample_idx, shuffle_idx


class GPTDataset(torch.utils.data.Dataset):
    def __init__(self, name: str, indexed: MMapIndexedDataset, documents: np.ndarray, seq_length: int,
                 num_samples: int, seed: int, cache_dir: str | Path | None = None):
        self.indexed, self.seq_length = indexed, seq_length
        key = hashlib.md5(f"{indexed.prefix}|{name}|{documents[0]}-{documents[-1]}|{seq_length}|"
                          f"{num_samples}|{seed}".encode()).hexdigest()[:12]
        cache = Path(cache_dir) if cache_dir else Path(indexed.prefix).parent / "index-cache"
        files = [cache / f"{name}_{key}_{k}.npy" for k in ("doc_idx", "sample_idx", "shuffle_idx")]
        if all(f.exists() for f in files):
            self.doc_idx, self.sample_idx, self.shuffle_idx = (np.load(f, mmap_mode="r") for f in files)
        else:
            self.doc_idx, self.sample_idx, self.shuffle_idx = build_indices(
                np.asarray(indexed.sizes), documents, seq_length, num_samples, seed)
            try:
                cache.mkdir(parents=True, exist_ok=True)
                for f, a in zip(files, (self.doc_idx, self.sample_idx, self.shuffle_idx)):
                    np.save(f, a)
            except OSError:
                pass

    def __len__(self) -> int:
        return len(self.shuffle_idx)

    def __getitem__(self, i: int) -> dict:
        s = int(self.shuffle_idx[i % len(self.shuffle_idx)])
        d0, o0 = (int(v) for v in self.sample_idx[s])
        d1, o1 = (int(v) for v in self.sample_idx[s + 1])
        if d0 == d1:
            toks = self.indexed.get(int(self.doc_idx[d0]), o0, o1 - o0 + 1)
        else:
            parts = [self.indexed.get(int(self.doc_idx[d0]), o0)]
            parts += [self.indexed.get(int(self.doc_idx[d])) for d in range(d0 + 1, d1)]
            parts.append(self.indexed.get(int(self.doc_idx[d1]), 0, o1 + 1))
            toks = np.concatenate(parts)
        return {"text": torch.from_numpy(np.asarray(toks, dtype=np.int64))}


def build_train_valid_test(prefix: str, split: str, seq_length: int, samples: tuple[int, int, int],
                           seed: int, cache_dir=None):
    ds = MMapIndexedDataset(prefix)
    ndocs = len(ds.doc_idx) - 1
    # sequences == documents for corpora written one item per document
    out = []
    for name, (lo, hi), n in zip(("train", "valid", "test"), parse_split(split, ndocs), samples):
        if hi - lo < 1 or n <= 0:
            out.append(None)
            continue
        docs = np.arange(lo, hi, dtype=np.int32)
        try:
            out.append(GPTDataset(name, ds, docs, seq_length, n, seed, cache_dir))
        except AssertionError:
            out.append(None)
    return out


def synthetic_corpus(prefix: str | Path, vocab_size: int, num_docs: int = 2000, mean_len: int = 600,
                     seed: int = 0) -> str:
    """Write a random-token corpus in the mmap format (used when no real corpus is mounted)."""
    from .indexed_dataset import MMapIndexedDatasetBuilder, best_dtype
    from .synthetic import token_stream

    b = MMapIndexedDatasetBuilder(prefix, best_dtype(vocab_size))
    for doc in token_stream(num_docs, vocab_size, mean_len, seed):
        b.add_item(doc)
        b.end_document()
    b.finalize()
    return str(prefix)
