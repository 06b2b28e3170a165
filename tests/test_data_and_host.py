"""Data layer and the C++ host library (libb200flow_host.so), on CPU.

* Megatron mmap `.bin/.idx` writer / reader round trip (megatron/data/indexed_dataset.py upstream;
  the reference prepares it with tools/preprocess_data.py,
  examples/megatron-deepspeed/gpt2_345m/wikicorpus.yaml:24-36).
* The C++ index builders (megatron/data/helpers.cpp upstream) against a plain Python walk.
* GPT-2 byte-level BPE on a locally written vocab / merges pair.
* Bucket planner (torch DDP Reducer bucket assignment) and the shared-memory mailbox.
"""
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


def _sentence_corpus(tmp_path, rng):
    from b200flow.data.indexed_dataset import MMapIndexedDatasetBuilder

    prefix = str(tmp_path / "bert_sentences")
    b = MMapIndexedDatasetBuilder(prefix, dtype=np.int32)
    doc_sents = []
    for d in range(60):
        nsent = int(rng.randint(1, 12))               # some single-sentence documents: skipped (min_num_sent = 2)
        lens = rng.randint(3, 40, size=nsent)
        if d == 7:
            lens[0] = 600                             # a sentence beyond the long-sentence limit: document skipped
        for n in lens:
            b.add_item(rng.randint(5, 1000, size=int(n)))
        b.end_document()
        doc_sents.append(lens.tolist())
    b.finalize()
    return prefix, doc_sents


def _python_bert_mapping(docs, sizes, num_epochs, max_num_samples, max_len, short_prob, seed, min_sent=2, long_len=512):
    """The same walk in Python, drawing from the same Mersenne-Twister stream."""
    rs = np.random.RandomState(seed)
    draw = lambda: int(rs.randint(0, 2 ** 32, dtype=np.uint64))     # noqa: E731
    every = int(1.0 / short_prob + 0.5) if short_prob > 0 else 0

    def target():
        r = draw()
        if every > 0 and r % every == 0:
            return 2 + draw() % (max_len - 1)
        return max_len

    rows = []
    for _ in range(num_epochs):
        for d in range(len(docs) - 1):
            first, last = int(docs[d]), int(docs[d + 1])
            if last - first < min_sent or any(sizes[s] > long_len for s in range(first, last)):
                continue
            start, tok, ns, tgt = first, 0, 0, target()
            for s in range(first, last):
                if len(rows) >= max_num_samples:
                    break
                tok += int(sizes[s])
                ns += 1
                rem = last - s - 1
                if (tok >= tgt and rem > 1 and ns >= min_sent) or rem == 0:
                    rows.append((start, s + 1, tgt))
                    start, tok, ns, tgt = s + 1, 0, 0, target()
    return rows[:max_num_samples]


def test_bert_sentence_mapping_cpp_matches_python_walk(tmp_path):
    from b200flow.data.bert_dataset import build_bert_mapping
    from b200flow.data.indexed_dataset import MMapIndexedDataset

    prefix, doc_sents = _sentence_corpus(tmp_path, np.random.RandomState(3))
    ds = MMapIndexedDataset(prefix)
    docs, sizes = np.asarray(ds.doc_idx), np.asarray(ds.sizes)
    assert len(docs) == 61 and len(sizes) == sum(len(x) for x in doc_sents)
    for epochs, cap, prob in ((1, 10 ** 9, 0.0), (3, 10 ** 9, 0.1), (5, 40, 0.25)):
        got = build_bert_mapping(docs, sizes, num_epochs=epochs, max_num_samples=cap, max_seq_length=61,
                                 short_seq_prob=prob, seed=99)
        want = _python_bert_mapping(docs, sizes, epochs, cap, 61, prob, 99)
        assert len(got) == len(want) and len(got) > 0
        assert sorted(map(tuple, got.tolist())) == sorted(want)            # same rows, shuffled order
        assert got.tolist() != sorted(got.tolist())
        again = build_bert_mapping(docs, sizes, num_epochs=epochs, max_num_samples=cap, max_seq_length=61,
                                   short_seq_prob=prob, seed=99)
        assert np.array_equal(got, again)                                   # deterministic
    # every row stays inside one document and skips document 7 (over-long sentence) and 1-sentence documents
    bounds = docs.tolist()
    for first, last, tgt in got.tolist():
        d = max(i for i in range(60) if bounds[i] <= first)
        assert last <= bounds[d + 1] and d != 7 and len(doc_sents[d]) >= 2 and 2 <= tgt <= 61


def test_bert_pretraining_samples(tmp_path):
    from b200flow.data.bert_dataset import BertPretrainingDataset

    prefix, _ = _sentence_corpus(tmp_path, np.random.RandomState(4))
    ds = BertPretrainingDataset(prefix, vocab_size=1000, cls_id=1, sep_id=2, mask_id=3, pad_id=0, max_seq_length=64,
                                num_epochs=2, seed=7)
    assert len(ds) > 50
    nsp = []
    for i in range(0, len(ds), 3):
        s = ds[i]
        ids, tt, am, lab = s["input_ids"], s["token_type_ids"], s["attention_mask"], s["labels"]
        n = int(am.sum())
        assert ids.shape == (64,) and n <= 64 and ids[0] == 1 and ids[n - 1] == 2 and (ids[n:] == 0).all()
        seps = (ids[:n] == 2).nonzero().flatten()
        assert len(seps) == 2 and (tt[:seps[0] + 1] == 0).all() and (tt[seps[0] + 1:n] == 1).all()
        pred = (lab != -100).nonzero().flatten()
        assert 1 <= len(pred) <= max(1, round(n * 0.15)) and (pred < n).all()
        assert not any(int(ids[p]) in (1, 2) and int(lab[p]) in (1, 2) for p in pred)   # specials are never predicted
        unchanged = (lab == -100)[:n]
        assert ((ids[:n] != 3) | ~unchanged).all()                                      # [MASK] only at predicted positions
        nsp.append(int(s["next_sentence_label"]))
        assert torch.equal(ds[i]["input_ids"], ids)                                     # a sample is a pure function of its index
    assert 0 < sum(nsp) < len(nsp)


def test_wordpiece_tokenizer_and_sentence_level_preprocessing(tmp_path):
    import json
    import subprocess
    import sys
    from pathlib import Path

    from b200flow.data.bert_dataset import BertPretrainingDataset
    from b200flow.data.tokenizer import BertWordPieceTokenizer, split_sentences, write_wordpiece_vocab

    vocab = tmp_path / "vocab.txt"
    vocab.write_text("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", "the", "un", "##aff", "##able", "runn",
                                "##ing", ",", ".", "he", "is", "cafe", "!"]) + "\n")
    tok = BertWordPieceTokenizer(vocab, lower_case=True)
    ids = tok.tokenize("He is UNAFFABLE, running. Café!  xyzzy")
    toks = [tok.inv[i] for i in ids]
    assert toks == ["he", "is", "un", "##aff", "##able", ",", "runn", "##ing", ".", "cafe", "!", "[UNK]"]
    assert tok.detokenize(ids[:5]) == "he is unaffable"
    pair, types = tok.encode_pair("he is", "the cafe", max_length=6)
    assert [tok.inv[i] for i in pair] == ["[CLS]", "he", "[SEP]", "the", "cafe", "[SEP]"] or len(pair) == 6
    assert len(pair) == len(types) == 6 and types[0] == 0 and types[-1] == 1
    assert BertWordPieceTokenizer(vocab, lower_case=False).tokenize("He") == [tok.unk]     # cased: "He" not in vocab
    assert split_sentences('It rained. "Why?" he asked (twice)! No answer came.\nNew line') == [
        "It rained.", '"Why?" he asked (twice)!', "No answer came.", "New line"]

    # synthetic vocabulary + sentence-level corpus through the Megatron preprocess tool, then BERT samples from it
    docs = [" ".join(f"Sentence number {i} of document {d} talks about topic {(i * d) % 7}." for i in range(2 + d % 5))
            for d in range(30)]
    big = tmp_path / "bert-vocab.txt"
    write_wordpiece_vocab(big, docs, size=400)
    src = tmp_path / "train.json"
    src.write_text("\n".join(json.dumps({"text": t}) for t in docs) + "\n")
    tool = Path(__file__).resolve().parent.parent / "b200flow/workloads/trees/megatron_deepspeed/tools/preprocess_data.py"
    res = subprocess.run([sys.executable, str(tool), "--input", str(src), "--output-prefix", str(tmp_path / "bert"),
                          "--vocab-file", str(big), "--tokenizer-type", "BertWordPieceLowerCase", "--split-sentences",
                          "--workers", "2"], capture_output=True, text=True, timeout=300,
                         env=dict(__import__("os").environ, PYTHONPATH=str(tool.parents[5])))
    assert res.returncode == 0, res.stdout + res.stderr
    t2 = BertWordPieceTokenizer(big)
    ds = BertPretrainingDataset(str(tmp_path / "bert_text_sentence"), vocab_size=t2.vocab_size, cls_id=t2.cls,
                                sep_id=t2.sep, mask_id=t2.mask, pad_id=t2.pad, max_seq_length=48, num_epochs=1)
    assert len(ds.ds.doc_idx) == 31 and len(ds.ds) == sum(2 + d % 5 for d in range(30)) and len(ds) >= 30
    s = ds[0]
    assert s["input_ids"][0] == t2.cls and t2.unk not in s["input_ids"].tolist()   # character fallback: nothing is unknown


class _SlowSquares:
    """Picklable toy dataset: worker 0's items are slow, so ordering cannot come from timing."""

    def __len__(self):
        return 64

    def __getitem__(self, i):
        import time

        if i == 13:
            raise ValueError("bad sample 13")
        if i % 2 == 0:
            time.sleep(0.01)
        return {"i": i, "x": torch.full((1000,), float(i * i))}


def test_prefetch_loader_keeps_order_and_reports_worker_errors():
    from b200flow.data.prefetch import PrefetchLoader

    order = [5, 2, 8, 1, 40, 41, 42, 3, 0, 63, 62, 7]
    got = list(PrefetchLoader(_SlowSquares(), order, workers=3, capacity=1 << 20))
    assert [g["i"] for g in got] == order and all(float(g["x"][0]) == g["i"] ** 2 for g in got)
    assert [g["i"] for g in PrefetchLoader(_SlowSquares(), order, workers=0)] == order
    with pytest.raises(RuntimeError, match="bad sample 13"):
        list(PrefetchLoader(_SlowSquares(), [1, 2, 13, 4], workers=2, capacity=1 << 20))


def test_blended_corpora_follow_their_weights(tmp_path):
    """`--data-path 0.25 A 0.75 B`: every prefix of the sample stream is mixed in proportion, each sample comes
    from the corpus the index names (corpus A only holds token 7, corpus B only token 9)."""
    from b200flow.data.gpt_dataset import BlendableDataset, build_train_valid_test_blended, parse_data_path

    for name, tok in (("a", 7), ("b", 9)):
        b = MMapIndexedDatasetBuilder(str(tmp_path / name), dtype=np.int32)
        for d in range(40):
            b.add_item(np.full(50 + d, tok, dtype=np.int32))
            b.end_document()
        b.finalize()
    assert parse_data_path(["0.25", "x", "0.75", "y"]) == ([0.25, 0.75], ["x", "y"])
    args = ["0.25", str(tmp_path / "a"), "0.75", str(tmp_path / "b")]
    train, valid, test = build_train_valid_test_blended(args, "90,10,0", 16, (200, 20, 0), seed=3,
                                                        cache_dir=tmp_path / "cache")
    assert isinstance(train, BlendableDataset) and len(train) == 200 and len(valid) == 20 and test is None
    toks = [int(train[i]["text"][0]) for i in range(200)]
    assert all((train[i]["text"] == toks[i]).all() and train[i]["text"].numel() == 17 for i in range(0, 200, 7))
    for n in (20, 80, 200):                          # proportion holds on every prefix, not only at the end
        share_a = sum(t == 7 for t in toks[:n]) / n
        assert abs(share_a - 0.25) <= 1.0 / n + 1e-9, (n, share_a)
    single = build_train_valid_test_blended([str(tmp_path / "a")], "90,10,0", 16, (50, 5, 0), seed=3)
    assert not isinstance(single[0], BlendableDataset) and len(single[0]) >= 50
