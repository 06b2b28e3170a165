"""GPT-2 byte-level BPE tokenizer hook (`--tokenizer-type GPT2BPETokenizer --vocab-file
gpt2-vocab.json --merge-file gpt2-merges.txt`, examples/megatron-deepspeed/gpt2_345m/
wikicorpus.yaml:36-41).  Reads the standard vocab.json / merges.txt pair when present; with no
network the launcher's vocab shim writes a small byte-level pair in the same format, and this
class then degenerates to byte tokens plus the learned merges."""
from __future__ import annotations

import json
from functools import lru_cache
from pathlib import Path

try:
    import regex as re
    _PAT = re.compile(r"""'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+""")
except ImportError:  # pragma: no cover
    import re
    _PAT = re.compile(r"""'s|'t|'re|'ve|'m|'ll|'d| ?[A-Za-z]+| ?[0-9]+| ?[^\sA-Za-z0-9]+|\s+(?!\S)|\s+""")


@lru_cache()
def bytes_to_unicode() -> dict[int, str]:
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    cs, n = bs[:], 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, (chr(c) for c in cs)))


class GPT2BPETokenizer:
    def __init__(self, vocab_file: str | Path, merge_file: str | Path):
        self.encoder = json.loads(Path(vocab_file).read_text())
        self.decoder = {v: k for k, v in self.encoder.items()}
        lines = Path(merge_file).read_text().split("\n")
        merges = [tuple(l.split()) for l in lines if l and not l.startswith("#") and len(l.split()) == 2]
        self.ranks = {m: i for i, m in enumerate(merges)}
        self.b2u = bytes_to_unicode()
        self.u2b = {v: k for k, v in self.b2u.items()}
        self.eod = self.encoder.get("<|endoftext|>", len(self.encoder) - 1)
        self._cache: dict[str, list[str]] = {}

    @property
    def vocab_size(self) -> int:
        return len(self.encoder)

    def _bpe(self, token: str) -> list[str]:
        if token in self._cache:
            return self._cache[token]
        word = list(token)
        while len(word) > 1:
            pairs = [(self.ranks.get((a, b), 1 << 30), i) for i, (a, b) in enumerate(zip(word, word[1:]))]
            rank, i = min(pairs)
            if rank == 1 << 30:
                break
            word[i:i + 2] = [word[i] + word[i + 1]]
        self._cache[token] = word
        return word

    def tokenize(self, text: str) -> list[int]:
        ids = []
        for tok in _PAT.findall(text):
            mapped = "".join(self.b2u[b] for b in tok.encode("utf-8"))
            for piece in self._bpe(mapped):
                if piece in self.encoder:
                    ids.append(self.encoder[piece])
                else:
                    ids.extend(self.encoder[c] for c in piece if c in self.encoder)
        return ids

    def detokenize(self, ids) -> str:
        text = "".join(self.decoder.get(int(i), "") for i in ids)
        return bytearray(self.u2b[c] for c in text if c in self.u2b).decode("utf-8", errors="replace")


def write_byte_level_vocab(vocab_file: str | Path, merge_file: str | Path, vocab_size: int = 50257) -> None:
    """A valid vocab.json / merges.txt pair: 256 byte tokens, common-pair merges, padding entries
    up to `vocab_size` - 1 and `<|endoftext|>` last (same ids layout as GPT-2's files)."""
    b2u = bytes_to_unicode()
    enc = {b2u[b]: i for i, b in enumerate(range(256))}
    merges = []
    sp = b2u[ord(" ")]
    for a in "etaoinshrdlu":
        merges.append((sp, a))
    for a, b in ["th", "he", "in", "er", "an", "re", "on", "at", "en", "nd", "ti", "es", "or", "te"]:
        merges.append((a, b))
    for a, b in merges:
        if a + b not in enc:
            enc[a + b] = len(enc)
    while len(enc) < vocab_size - 1:
        enc[f"<|unused{len(enc)}|>"] = len(enc)
    enc["<|endoftext|>"] = len(enc)
    Path(vocab_file).write_text(json.dumps(enc))
    Path(merge_file).write_text("#version: 0.2\n" + "\n".join(f"{a} {b}" for a, b in merges) + "\n")


# ------------------------------------------------------------------ WordPiece (BERT)
def _is_punct(ch: str) -> bool:
    import unicodedata

    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


class BertWordPieceTokenizer:
    """BERT's tokenizer from a local `vocab.txt` (one token per line) — Megatron's
    `BertWordPieceLowerCase` / `BertWordPieceCase` tokenizer types and what `bert-base-cased` of the
    reference's Accelerate / Ray examples uses (examples/accelerate/bert-glue-mrpc/pretrain.yaml:45-50).

    Basic tokenisation (whitespace clean-up, optional lower-casing + accent stripping, punctuation and
    CJK characters as their own tokens) followed by greedy longest-match-first WordPiece with the
    `##` continuation prefix; words longer than 100 characters or without a match become `[UNK]`."""

    def __init__(self, vocab_file: str | Path, lower_case: bool = True, max_chars_per_word: int = 100):
        with open(vocab_file, encoding="utf-8") as f:
            toks = [l.rstrip("\n") for l in f]
        self.vocab = {t: i for i, t in enumerate(toks) if t != ""}
        self.inv = {i: t for t, i in self.vocab.items()}
        self.lower, self.max_chars = lower_case, max_chars_per_word
        self.unk, self.cls, self.sep = self.vocab["[UNK]"], self.vocab["[CLS]"], self.vocab["[SEP]"]
        self.mask, self.pad = self.vocab["[MASK]"], self.vocab["[PAD]"]
        self.eod = self.sep

    @property
    def vocab_size(self) -> int:
        return len(self.vocab)

    def _basic(self, text: str) -> list[str]:
        import unicodedata

        out, word = [], []

        def flush():
            if word:
                out.append("".join(word))
                word.clear()

        for ch in text:
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or (unicodedata.category(ch) in ("Cc", "Cf") and ch not in "\t\n\r"):
                continue
            if ch.isspace():
                flush()
            elif _is_punct(ch) or 0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0xF900 <= cp <= 0xFAFF:
                flush()
                out.append(ch)
            else:
                word.append(ch)
        flush()
        if self.lower:
            out = ["".join(c for c in unicodedata.normalize("NFD", w.lower()) if unicodedata.category(c) != "Mn")
                   for w in out]
        return [w for w in out if w]

    def _wordpiece(self, word: str) -> list[int]:
        if len(word) > self.max_chars:
            return [self.unk]
        ids, start = [], 0
        while start < len(word):
            end, cur = len(word), None
            while start < end:
                piece = ("##" if start > 0 else "") + word[start:end]
                if piece in self.vocab:
                    cur = self.vocab[piece]
                    break
                end -= 1
            if cur is None:
                return [self.unk]
            ids.append(cur)
            start = end
        return ids

    def tokenize(self, text: str) -> list[int]:
        return [i for w in self._basic(text) for i in self._wordpiece(w)]

    def encode_pair(self, a: str, b: str | None = None, max_length: int | None = None):
        """([CLS] a [SEP] b [SEP], token types) — longest-first truncation like HF's `truncation=True`."""
        ia, ib = self.tokenize(a), (self.tokenize(b) if b is not None else [])
        if max_length is not None:
            budget = max_length - (3 if b is not None else 2)
            while len(ia) + len(ib) > budget:
                (ia if len(ia) >= len(ib) else ib).pop()
        ids = [self.cls] + ia + [self.sep] + (ib + [self.sep] if b is not None else [])
        types = [0] * (len(ia) + 2) + [1] * (len(ib) + 1 if b is not None else 0)
        return ids, types

    def detokenize(self, ids) -> str:
        words: list[str] = []
        for i in ids:
            t = self.inv.get(int(i), "[UNK]")
            if t.startswith("##") and words:
                words[-1] += t[2:]
            else:
                words.append(t)
        return " ".join(words)


def split_sentences(text: str) -> list[str]:
    """Sentence boundaries for `preprocess_data.py --split-sentences` (Megatron uses nltk's punkt model, which
    is not in the image): a sentence ends at `.`, `!` or `?` (plus closing quotes / brackets) followed by
    whitespace and an upper-case letter, digit or opening quote; line breaks always end one."""
    import re

    out = []
    for para in text.splitlines():
        para = para.strip()
        if para:
            out += [s.strip() for s in re.split(r"(?<=[.!?])[\"')\]]*\s+(?=[\"'(\[]?[A-Z0-9])", para) if s.strip()]
    return out


def write_wordpiece_vocab(vocab_file: str | Path, texts=(), size: int = 2000, lower_case: bool = True) -> None:
    """A usable `vocab.txt` without network access: the special tokens, every printable ASCII character and
    its `##` form (so nothing is ever out of vocabulary), then the most frequent words and word-final /
    word-internal fragments of `texts`."""
    from collections import Counter

    specials = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    chars = [chr(c) for c in range(33, 127)]
    if lower_case:
        chars = [c for c in chars if not ("A" <= c <= "Z")]
    vocab = specials + chars + ["##" + c for c in chars]
    seen = set(vocab)
    counts: Counter = Counter()
    for t in texts:
        for w in (t.lower() if lower_case else t).split():
            w = "".join(c for c in w if c.isalnum())
            if not w:
                continue
            counts[w] += 3
            for k in range(2, min(len(w), 6)):
                counts["##" + w[-k:]] += 1
                counts[w[:k]] += 1
    for tok, _ in counts.most_common():
        if len(vocab) >= size:
            break
        if tok not in seen:
            seen.add(tok)
            vocab.append(tok)
    Path(vocab_file).write_text("\n".join(vocab) + "\n", encoding="utf-8")
