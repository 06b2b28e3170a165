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
