"""Offline stand-in for the HuggingFace `datasets` entry point the reference's scripts call with a
hub name (`load_dataset("wikicorpus", "raw_en", split="train")`,
examples/megatron-deepspeed/gpt2_345m/wikicorpus.yaml:15-22; `load_dataset("glue", "cola")` in
the Ray example).  The launcher puts this directory first on PYTHONPATH only in offline mode;
it returns synthetic data of the named shape.  `B2F_SYNTH_DOCS` sizes text corpora."""
import json
import os

from b200flow.data import synthetic as _syn


class _TextDataset:
    def __init__(self, rows):
        self.rows = rows
        self.column_names = list(rows[0].keys()) if rows else []

    def __len__(self):
        return len(self.rows)

    def __iter__(self):
        return iter(self.rows)

    def __getitem__(self, i):
        return self.rows[i]

    def to_json(self, path, **kw):
        with open(path, "w", encoding="utf-8") as f:
            for r in self.rows:
                f.write(json.dumps(r) + "\n")
        return os.path.getsize(path)

    def select(self, idx):
        return _TextDataset([self.rows[i] for i in idx])


def load_dataset(path, name=None, split=None, **kw):
    key = (str(path), str(name))
    if path == "wikicorpus":
        n = int(os.environ.get("B2F_SYNTH_DOCS", 2000))
        return _TextDataset(list(_syn.text_documents(n)))
    if path == "glue":
        maker = {"mrpc": _syn.mrpc_like, "cola": _syn.cola_like}[name]
        splits = {s: _TextDataset(maker(s)) for s in ("train", "validation")}
        return splits[split] if split else splits
    raise FileNotFoundError(f"[b200flow offline] no synthetic stand-in for dataset {key}")
