"""Memory-mapped indexed token dataset (`.bin` + `.idx`), the on-disk format Megatron's
`tools/preprocess_data.py --dataset-impl mmap` produces and `--data-impl mmap` consumes
(examples/megatron-deepspeed/gpt2_345m/wikicorpus.yaml:34-42 writes `gpt2_text_document.{bin,idx}`;
pretrain-ddp-zero1.yaml:28,54 reads it).

`.idx` layout: magic `MMIDIDX\\0\\0`, u64 version (1), u8 dtype code, u64 #sequences,
u64 #documents, i32 sizes[#seq], i64 byte pointers[#seq], i64 doc_idx[#doc].  `.bin` is the
concatenated token arrays.
"""
from __future__ import annotations

import struct
from pathlib import Path

import numpy as np

_MAGIC = b"MMIDIDX\x00\x00"
_DTYPES = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float32,
           7: np.float64, 8: np.uint16}
_CODES = {np.dtype(v): k for k, v in _DTYPES.items()}


def best_dtype(vocab_size: int):
    return np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32


class MMapIndexedDatasetBuilder:
    def __init__(self, out_prefix: str | Path, dtype=np.int32):
        self.prefix = Path(str(out_prefix))
        self.dtype = np.dtype(dtype)
        self.prefix.parent.mkdir(parents=True, exist_ok=True)
        self._bin = open(str(self.prefix) + ".bin", "wb")
        self.sizes: list[int] = []
        self.doc_idx: list[int] = [0]

    def add_item(self, tokens) -> None:
        arr = np.asarray(tokens, dtype=self.dtype)
        self._bin.write(arr.tobytes(order="C"))
        self.sizes.append(arr.size)

    def end_document(self) -> None:
        self.doc_idx.append(len(self.sizes))

    def finalize(self) -> None:
        self._bin.close()
        sizes = np.asarray(self.sizes, dtype=np.int32)
        ptrs = np.zeros(len(sizes), dtype=np.int64)
        if len(sizes) > 1:
            np.cumsum(sizes[:-1].astype(np.int64) * self.dtype.itemsize, out=ptrs[1:])
        with open(str(self.prefix) + ".idx", "wb") as f:
            f.write(_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", _CODES[self.dtype]))
            f.write(struct.pack("<Q", len(sizes)))
            f.write(struct.pack("<Q", len(self.doc_idx)))
            f.write(sizes.tobytes(order="C"))
            f.write(ptrs.tobytes(order="C"))
            f.write(np.asarray(self.doc_idx, dtype=np.int64).tobytes(order="C"))


class MMapIndexedDataset:
    def __init__(self, prefix: str | Path):
        self.prefix = str(prefix)
        with open(self.prefix + ".idx", "rb") as f:
            assert f.read(9) == _MAGIC, "not an mmap indexed dataset index"
            (version,) = struct.unpack("<Q", f.read(8))
            assert version == 1
            (code,) = struct.unpack("<B", f.read(1))
            self.dtype = np.dtype(_DTYPES[code])
            (n,) = struct.unpack("<Q", f.read(8))
            (nd,) = struct.unpack("<Q", f.read(8))
            off = f.tell()
        idx = np.memmap(self.prefix + ".idx", mode="r", order="C")
        self.sizes = np.frombuffer(idx, dtype=np.int32, count=n, offset=off)
        self.pointers = np.frombuffer(idx, dtype=np.int64, count=n, offset=off + self.sizes.nbytes)
        self.doc_idx = np.frombuffer(idx, dtype=np.int64, count=nd,
                                     offset=off + self.sizes.nbytes + self.pointers.nbytes)
        self._bin = np.memmap(self.prefix + ".bin", mode="r", order="C")

    def __len__(self) -> int:
        return len(self.sizes)

    def get(self, i: int, offset: int = 0, length: int | None = None) -> np.ndarray:
        size = int(self.sizes[i])
        length = size - offset if length is None else length
        start = int(self.pointers[i]) + offset * self.dtype.itemsize
        return np.frombuffer(self._bin, dtype=self.dtype, count=length, offset=start)

    def __getitem__(self, i: int) -> np.ndarray:
        return self.get(i)

    @staticmethod
    def exists(prefix: str | Path) -> bool:
        return Path(str(prefix) + ".idx").exists() and Path(str(prefix) + ".bin").exists()
