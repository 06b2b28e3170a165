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
                # Several ranks may build the same arrays at once, and others memory-map the files as soon as all
                # three exist: write under a private name and rename, so a file that exists is complete (a
                # half-written one under a reader's mmap is a bus error).
                import os

                cache.mkdir(parents=True, exist_ok=True)
                for f, a in zip(files, (self.doc_idx, self.sample_idx, self.shuffle_idx)):
                    tmp = f.with_name(f"{f.stem}.{os.getpid()}.tmp.npy")
                    np.save(tmp, a)
                    os.replace(tmp, f)
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


class BlendableDataset(torch.utils.data.Dataset):
    """Megatron's `BlendableDataset`: samples drawn from several datasets in proportion to their weights
    (`--data-path 0.3 corpusA 0.7 corpusB`).  The assignment of every global sample to (dataset, sample within
    it) is built once by `b2f_build_blending_indices` (csrc/host/data_helpers.cpp): always the dataset whose
    realised share lags its target share the most, so any prefix of the stream is mixed in proportion."""

    def __init__(self, datasets: list, weights: list[float], size: int):
        assert len(datasets) == len(weights) and 0 < len(datasets) < 255
        w = np.asarray(weights, dtype=np.float64)
        assert (w > 0).all(), "blend weights must be positive"
        self.datasets, self.weights, self.size = datasets, w / w.sum(), int(size)
        self.dataset_index = np.zeros(self.size, dtype=np.uint8)
        self.dataset_sample_index = np.zeros(self.size, dtype=np.int64)
        _host().b2f_build_blending_indices(
            self.dataset_index.ctypes.data_as(ctypes.c_void_p), self.dataset_sample_index.ctypes.data_as(ctypes.c_void_p),
            self.weights.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(datasets)), ctypes.c_longlong(self.size))

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, i: int) -> dict:
        d = int(self.dataset_index[i % self.size])
        return self.datasets[d][int(self.dataset_sample_index[i % self.size])]


def parse_data_path(data_path: list[str]) -> tuple[list[float], list[str]]:
    """`--data-path` is either one prefix or `weight prefix weight prefix ...`."""
    if len(data_path) == 1:
        return [1.0], [data_path[0]]
    assert len(data_path) % 2 == 0, "--data-path takes one prefix or weight / prefix pairs"
    return [float(x) for x in data_path[0::2]], list(data_path[1::2])


def build_train_valid_test_blended(data_path: list[str], split: str, seq_length: int, samples: tuple[int, int, int],
                                   seed: int, cache_dir=None):
    """Train / valid / test datasets over one corpus or a weighted blend of corpora (Megatron
    `build_train_valid_test_datasets`): every corpus is split by `--split` on its own and contributes
    ceil(weight x samples x 1.005) samples to each blend, like upstream's safety margin."""
    weights, prefixes = parse_data_path(data_path)
    if len(prefixes) == 1:
        return build_train_valid_test(prefixes[0], split, seq_length, samples, seed, cache_dir)
    w = np.asarray(weights, dtype=np.float64)
    w = w / w.sum()
    per = [build_train_valid_test(p, split, seq_length,
                                  tuple(int(np.ceil(n * wi * 1.005)) for n in samples), seed, cache_dir)
           for p, wi in zip(prefixes, w)]
    out = []
    for k, n in enumerate(samples):
        parts = [(ds[k], wi) for ds, wi in zip(per, w) if ds[k] is not None]
        out.append(BlendableDataset([p for p, _ in parts], [wi for _, wi in parts], n) if parts and n > 0 else None)
    return out
