"""Data layer and the C++ host library (libb200flow_host.so), on CPU.

* Megatron mmap `.bin/.idx` writer / reader round trip (megatron/data/indexed_dataset.py upstream;
  the reference prepares it with tools/preprocess_data.py,
  examples/megatron-deepspeed/gpt2_345m/wikicorpus.yaml:24-36).
* The C++ index builders (megatron/data/helpers.cpp upstream) against a plain Python walk.
* GPT-2 byte-level BPE on a locally written vocab / merges pair.
* Bucket planner (torch DDP Reducer bucket assignment) and the shared-memory mailbox.
"""
import ctypes
import multiprocessing as mp

import numpy as np
import pytest
import torch

from b200flow.data.gpt_dataset import build_indices, build_train_valid_test, parse_split, synthetic_corpus
from b200flow.data.indexed_dataset import MMapIndexedDataset, MMapIndexedDatasetBuilder, best_dtype
from b200flow.data.tokenizer import GPT2BPETokenizer, write_byte_level_vocab


def test_mmap_indexed_dataset_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    docs = [rng.integers(0, 50000, size=int(n)).astype(np.int64) for n in rng.integers(1, 400, size=50)]
    prefix = tmp_path / "corpus_text_document"
    b = MMapIndexedDatasetBuilder(prefix, best_dtype(50257))
    for d in docs:
        b.add_item(d)
        b.end_document()
    b.finalize()
    assert MMapIndexedDataset.exists(prefix)
    ds = MMapIndexedDataset(prefix)
    assert len(ds) == len(docs)
    for i in (0, 7, 49):
        assert np.array_equal(ds[i].astype(np.int64), docs[i])
    long_doc = max(range(len(docs)), key=lambda i: len(docs[i]))
    assert np.array_equal(ds.get(long_doc, offset=3, length=5).astype(np.int64), docs[long_doc][3:8])
    assert best_dtype(50257) == np.uint16 and best_dtype(70000) == np.int32


def test_cpp_sample_index_matches_python_walk():
    rng = np.random.default_rng(1)
    sizes = rng.integers(5, 90, size=40).astype(np.int32)
    documents = np.arange(40, dtype=np.int32)
    seq = 32
    doc_idx, sample_idx, shuffle_idx = build_indices(sizes, documents, seq, num_samples=100, seed=1234)
    # doc_idx: whole epochs, each a permutation of the documents
    per = len(documents)
    assert len(doc_idx) % per == 0
    for e in range(len(doc_idx) // per):
        assert sorted(doc_idx[e * per:(e + 1) * per].tolist()) == list(range(per))
    # sample_idx: walking seq+1 tokens per sample (the last token is shared with the next sample)
    d, off = 0, 0
    assert sample_idx[0].tolist() == [0, 0]
    for s in range(1, len(sample_idx)):
        remaining = seq
        while remaining > 0:
            left = int(sizes[doc_idx[d]]) - off
            if left > remaining:
                off += remaining
                remaining = 0
            else:
                remaining -= left
                d, off = d + 1, 0
        assert sample_idx[s].tolist() == [d, off], s
    assert sorted(shuffle_idx.tolist()) == list(range(len(sample_idx) - 1))
    # same seed, same indices; another seed, another order
    again = build_indices(sizes, documents, seq, num_samples=100, seed=1234)
    assert all(np.array_equal(a, b) for a, b in zip((doc_idx, sample_idx, shuffle_idx), again))
    other = build_indices(sizes, documents, seq, num_samples=100, seed=99)
    assert not np.array_equal(other[2], shuffle_idx)


def test_gpt_dataset_splits_and_sample_shape(tmp_path):
    prefix = synthetic_corpus(tmp_path / "synth_text_document", vocab_size=1000, num_docs=200, mean_len=80, seed=3)
    assert parse_split("949,50,1", 1000) == [(0, 949), (949, 999), (999, 1000)]
    train, valid, test = build_train_valid_test(prefix, "8,1,1", seq_length=64, samples=(50, 10, 5), seed=7)
    assert train is not None and valid is not None
    assert len(train) >= 50
    item = train[0]["text"]
    assert item.dtype == torch.int64 and item.shape == (65,)      # seq_length + 1: inputs and shifted labels
    assert int(item.max()) < 1000 and int(item.min()) >= 0
    assert torch.equal(train[0]["text"], train[0]["text"])
    assert not torch.equal(train[0]["text"], train[1]["text"])


def test_gpt2_bpe_tokenizer_round_trip(tmp_path):
    vocab, merges = tmp_path / "gpt2-vocab.json", tmp_path / "gpt2-merges.txt"
    write_byte_level_vocab(vocab, merges, vocab_size=50257)
    tok = GPT2BPETokenizer(vocab, merges)
    assert tok.vocab_size == 50257
    text = "the other thing in there, then another: naïve café 123!"
    ids = tok.tokenize(text)
    assert all(0 <= i < 50257 for i in ids)
    assert tok.detokenize(ids) == text
    assert len(ids) < len(text.encode())       # some merges applied ("th", "he", " t", ...)


def test_bucket_planner_covers_all_slots_in_backward_order():
    from b200flow.parallel.ddp import plan_buckets
    from b200flow.utils.flat import FlatParams

    model = torch.nn.Sequential(*[torch.nn.Linear(256, 256) for _ in range(12)])
    flat = FlatParams(model, pad_to_multiple=128)
    buckets = plan_buckets(flat, cap_mb=0.5, first_cap_mb=0.1)
    n = len(flat.slots)
    covered = []
    for s0, s1, e0, e1 in buckets:
        assert 0 <= s0 < s1 <= n and e0 < e1
        assert e0 == flat.slots[s0].offset
        covered += list(range(s0, s1))
    assert sorted(covered) == list(range(n)) and len(set(covered)) == n
    # backward produces the last layers' gradients first: buckets come in reverse slot order
    firsts = [b[0] for b in buckets]
    assert firsts == sorted(firsts, reverse=True)
    # the first bucket is the small one (lets the first all-reduce start early), the rest obey the cap
    sizes = [(e1 - e0) * 4 for _, _, e0, e1 in buckets]
    assert sizes[0] <= 0.1 * (1 << 20) + 256 * 256 * 4 + 1024
    assert len(buckets) > 3


def _mailbox_child(name):
    from b200flow.engines.ray_train_api import Channel

    ch = Channel(name, create=False)
    for i in range(3):
        ch.send({"rank": 0, "step": i, "payload": list(range(i * 1000))})
    ch.send({"big": b"x" * (3 << 20)})      # larger than the receiver's first buffer
    ch.close()


def test_shared_memory_mailbox_between_processes():
    import os

    from b200flow.engines.ray_train_api import Channel

    name = f"/b2f_test_{os.getpid()}"
    ch = Channel(name, create=True, capacity=8 << 20)
    try:
        p = mp.get_context("spawn").Process(target=_mailbox_child, args=(name,))
        p.start()
        got = []
        for _ in range(600):
            m = ch.recv(timeout_ms=100)
            if m is not None:
                got.append(m)
            if len(got) == 4:
                break
        p.join(timeout=30)
        assert p.exitcode == 0
        assert [m["step"] for m in got[:3]] == [0, 1, 2]
        assert got[2]["payload"] == list(range(2000))
        assert len(got[3]["big"]) == 3 << 20
        with pytest.raises(EOFError):          # drained and the writer closed its end
            ch.recv(timeout_ms=10)
    finally:
        ch.free()


def test_polygon_rasteriser_oracle_and_rle_round_trip():
    from b200flow.ops.detection import poly_to_mask_ref, rle_area, rle_decode, rle_encode, rle_iou

    objects = [[[2, 2, 12, 2, 12, 9, 2, 9]],                              # axis-aligned box
               [[1, 1, 6, 1, 6, 6, 1, 6], [10, 10, 15, 10, 15, 15, 10, 15]],   # two parts
               [[8.0, 1.0, 15.0, 15.0, 1.0, 15.0]]]                         # triangle
    m = poly_to_mask_ref(objects, 16, 16)
    assert m.shape == (3, 16, 16) and m.dtype == torch.uint8
    assert int(m[0].sum()) == 10 * 7 and bool(m[0, 2, 2]) and not bool(m[0, 9, 2])
    assert int(m[1].sum()) == 25 + 25
    assert 0 < int(m[2].sum()) < 14 * 14 // 2 + 14
    for k in range(3):
        r = rle_encode(m[k])
        assert r["size"] == [16, 16] and sum(r["counts"]) == 256
        assert torch.equal(rle_decode(r), m[k]) and rle_area(r) == int(m[k].sum())
    full = rle_encode(torch.ones(4, 5, dtype=torch.uint8))
    assert full["counts"] == [0, 20]
    assert rle_iou(rle_encode(m[0]), rle_encode(m[0])) == 1.0
    assert 0.0 < rle_iou(rle_encode(m[0]), rle_encode(m[2])) < 1.0
