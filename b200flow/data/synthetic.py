"""Synthetic data of the shapes the reference workloads train on (no network on the box).

* `mrpc_like` / `cola_like` : GLUE-shaped sentence(-pair) classification sets — sizes of the
  real ones (MRPC 3668/408, CoLA 8551/1043), variable lengths, a learnable label signal
  (the label decides whether a marker token is planted), so accuracy/F1/MCC move with
  training like the reference's in-workload evaluation.
* `token_stream`            : documents for the GPT preprocessing / mmap pipeline.
* `coco_like`               : COCO-2017-shaped detection samples (images, boxes, class ids,
  instance masks) generated on the fly; `write_coco_tree` lays out the directory the
  `coco-data` chart creates (charts/machine-learning/data-prep/coco-data/templates/coco-data.yaml:13-35).
* `python -m b200flow.data.synthetic coco|redpajama --out DIR` is what the data charts run.
"""
from __future__ import annotations

import argparse
import json
import math
from pathlib import Path

import numpy as np
import torch

CLS, SEP, PAD, MARK_A, MARK_B = 101, 102, 0, 2000, 2001


def _glue_like(n: int, pair: bool, vocab: int, max_len: int, seed: int):
    rng = np.random.default_rng(seed)
    rows = []
    lo = min(1000, vocab // 2)
    mark_a, mark_b = min(MARK_A, vocab - 2), min(MARK_B, vocab - 1)
    hi = min(vocab, mark_a) if mark_a < 2000 else vocab
    for _ in range(n):
        label = int(rng.integers(0, 2))
        la = int(rng.integers(5, max_len // 2 - 2))
        a = rng.integers(lo, hi, la).tolist()
        if label:
            a[int(rng.integers(0, la))] = mark_a
        ids = [CLS] + a + [SEP]
        types = [0] * len(ids)
        if pair:
            lb = int(rng.integers(5, max_len // 2 - 2))
            b = rng.integers(lo, hi, lb).tolist()
            if label:
                b[int(rng.integers(0, lb))] = mark_b
            ids += b + [SEP]
            types += [1] * (lb + 1)
        rows.append({"input_ids": ids, "token_type_ids": types, "labels": label})
    return rows


def mrpc_like(split: str = "train", vocab: int = 28996, max_len: int = 128, seed: int = 42):
    n = {"train": 3668, "validation": 408, "test": 1725}[split]
    return _glue_like(n, True, vocab, max_len, seed + {"train": 0, "validation": 1, "test": 2}[split])


def cola_like(split: str = "train", vocab: int = 28996, max_len: int = 64, seed: int = 7):
    n = {"train": 8551, "validation": 1043}[split]
    return _glue_like(n, False, vocab, max_len, seed + (0 if split == "train" else 1))


def collate_pad(rows, pad_to_multiple_of: int = 8):
    """Pad to the longest sequence in the batch, rounded up (the example's collate_fn)."""
    m = max(len(r["input_ids"]) for r in rows)
    m = (m + pad_to_multiple_of - 1) // pad_to_multiple_of * pad_to_multiple_of
    ids = torch.full((len(rows), m), PAD, dtype=torch.long)
    types = torch.zeros((len(rows), m), dtype=torch.long)
    mask = torch.zeros((len(rows), m), dtype=torch.long)
    for i, r in enumerate(rows):
        n = len(r["input_ids"])
        ids[i, :n] = torch.tensor(r["input_ids"])
        types[i, :n] = torch.tensor(r["token_type_ids"])
        mask[i, :n] = 1
    return {"input_ids": ids, "token_type_ids": types, "attention_mask": mask,
            "labels": torch.tensor([r["labels"] for r in rows])}


def token_stream(num_docs: int, vocab: int = 50257, mean_len: int = 400, seed: int = 0):
    rng = np.random.default_rng(seed)
    for _ in range(num_docs):
        n = max(8, int(rng.normal(mean_len, mean_len / 4)))
        yield rng.integers(0, vocab, n, dtype=np.int64)


_ZIPF_CDF: dict = {}


def zipf_tokens(vocab: int, shape, generator=None, exponent: float = 1.1, offset: float = 2.7):
    """Token ids with a natural-language-like (Zipf-Mandelbrot) unigram distribution, p(r) ~ 1 / (r + offset)^exponent
    over a fixed pseudo-random assignment of ranks to ids.  The benches train on these instead of uniform noise so
    that a loss curve means something: a language model first learns the unigram distribution, and the loss falls from
    ~ln(vocab) + initial logit noise towards the entropy of the distribution (about 7 nats at vocab 50k) - or it does
    not, which is what a broken fp8 path or optimizer would show."""
    key = (vocab, exponent, offset)
    if key not in _ZIPF_CDF:
        r = torch.arange(vocab, dtype=torch.float64)
        w = (r + offset).pow(-exponent)
        perm = torch.randperm(vocab, generator=torch.Generator().manual_seed(97))
        _ZIPF_CDF[key] = ((w / w.sum()).cumsum(0), perm)
    cdf, perm = _ZIPF_CDF[key]
    u = torch.rand(shape, generator=generator, dtype=torch.float64)
    return perm[torch.searchsorted(cdf, u).clamp_(max=vocab - 1)]


def text_documents(num_docs: int, seed: int = 0):
    """Plain-text pseudo documents (json lines with a "text" field, as wikicorpus -> train.json)."""
    rng = np.random.default_rng(seed)
    words = ["alpha", "beta", "gamma", "delta", "epsilon", "zeta", "eta", "theta", "iota", "kappa",
             "lambda", "mu", "nu", "xi", "omicron", "pi", "rho", "sigma", "tau", "upsilon"]
    for _ in range(num_docs):
        n = int(rng.integers(20, 200))
        yield {"text": " ".join(words[int(i)] for i in rng.integers(0, len(words), n)) + "."}


def coco_like(num_images: int, height: int = 800, width: int = 1344, max_objects: int = 12,
              num_classes: int = 80, seed: int = 0, device="cpu", mask: bool = True):
    """Yields dicts: image [3,H,W] float, boxes [n,4] xyxy, labels [n] in 1..80, masks [n,H,W]."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    for _ in range(num_images):
        n = int(torch.randint(1, max_objects + 1, (1,), generator=g))
        cx = torch.rand(n, generator=g) * width
        cy = torch.rand(n, generator=g) * height
        w = (torch.rand(n, generator=g) * 0.4 + 0.05) * width
        h = (torch.rand(n, generator=g) * 0.4 + 0.05) * height
        boxes = torch.stack([(cx - w / 2).clamp(0, width - 2), (cy - h / 2).clamp(0, height - 2),
                             (cx + w / 2).clamp(1, width - 1), (cy + h / 2).clamp(1, height - 1)], 1)
        boxes[:, 2:] = torch.maximum(boxes[:, 2:], boxes[:, :2] + 4)
        labels = torch.randint(1, num_classes + 1, (n,), generator=g)
        img = torch.rand(3, height, width, generator=g) * 0.2
        masks = torch.zeros(n, height, width, dtype=torch.uint8) if mask else None
        ys, xs = torch.arange(height)[:, None], torch.arange(width)[None, :]
        for i in range(n):
            x0, y0, x1, y1 = boxes[i].tolist()
            ell = (((xs - (x0 + x1) / 2) / max((x1 - x0) / 2, 1)) ** 2 +
                   ((ys - (y0 + y1) / 2) / max((y1 - y0) / 2, 1)) ** 2) <= 1.0
            if mask:
                masks[i] = ell
            img[:, ell] += (labels[i].float() / num_classes) * 0.5 + 0.2   # class-dependent brightness
        out = {"image": img.to(device), "boxes": boxes.to(device), "labels": labels.to(device)}
        if mask:
            out["masks"] = masks.to(device)
        yield out


COCO_SPLITS = {"train2017": 16, "val2017": 4, "test2017": 4}
from ..launcher.offline import STANDIN_KEY  # noqa: E402  (the marker the offline wget / unzip shims exchange)


def _coco_split(split: str):
    """The synthetic samples of one COCO split; a function of the split name only, so images and annotations can be
    written by different processes (the download chart unpacks them from different archives)."""
    import zlib

    return coco_like(COCO_SPLITS[split], 128, 160, seed=zlib.crc32(split.encode()) % 1000, mask=False)


def write_coco_images(out: Path, split: str) -> None:
    (out / split).mkdir(parents=True, exist_ok=True)
    for i, s in enumerate(_coco_split(split)):
        torch.save(s["image"].half(), out / split / f"{i:012d}.pt")


def write_coco_annotations(out: Path, splits=("train2017", "val2017")) -> None:
    (out / "annotations").mkdir(parents=True, exist_ok=True)
    for split in splits:
        images, anns = [], []
        for i, s in enumerate(_coco_split(split)):
            images.append({"id": i, "file_name": f"{i:012d}.pt", "height": 128, "width": 160})
            for b, l in zip(s["boxes"].tolist(), s["labels"].tolist()):
                cx, cy, rx, ry = (b[0] + b[2]) / 2, (b[1] + b[3]) / 2, (b[2] - b[0]) / 2, (b[3] - b[1]) / 2
                poly = [v for k in range(16) for v in (cx + rx * math.cos(2 * math.pi * k / 16),
                                                       cy + ry * math.sin(2 * math.pi * k / 16))]
                anns.append({"image_id": i, "bbox": [b[0], b[1], b[2] - b[0], b[3] - b[1]],
                             "category_id": l, "iscrowd": 0, "id": len(anns), "segmentation": [poly],
                             "area": math.pi * rx * ry})
        (out / "annotations" / f"instances_{split}.json").write_text(json.dumps(
            {"images": images, "annotations": anns,
             "categories": [{"id": c, "name": f"class{c}"} for c in range(1, 81)]}))


def write_backbone_placeholder(path: Path) -> None:
    path.parent.mkdir(parents=True, exist_ok=True)
    np.savez(path, synthetic=np.zeros(1))


def write_coco_tree(out: Path, num_train: int = 16, num_val: int = 4) -> None:
    """Directory layout of the coco-data chart: train2017/ val2017/ annotations/ pretrained-models/."""
    out = Path(out)
    for split in ("train2017", "val2017"):
        write_coco_images(out, split)
    write_coco_annotations(out)
    write_backbone_placeholder(out / "pretrained-models" / "ImageNet-R50-AlignPadding.npz")


def write_standin(url: str, dst: Path, docs: int = 2000) -> str:
    """What the offline `wget` shim leaves in place of a download (launcher/offline.py): data files the workloads read
    directly are generated at once (`.npz` backbone placeholder, `.jsonl` text corpus); an archive becomes a small marker
    that the `unzip` shim expands with `unpack_standin`."""
    dst = Path(dst)
    dst.parent.mkdir(parents=True, exist_ok=True)
    name = url.rstrip("/").rsplit("/", 1)[-1]
    if name.endswith(".npz"):
        write_backbone_placeholder(dst)
        return "backbone placeholder"
    if name.endswith(".jsonl"):
        with open(dst, "w") as f:
            for d in text_documents(docs):
                f.write(json.dumps(d) + "\n")
        return f"{docs} synthetic documents"
    dst.write_text(json.dumps({STANDIN_KEY: url, "name": name}) + "\n")
    return "archive marker"


def read_standin(archive: Path) -> dict | None:
    try:
        with open(archive, "rb") as f:
            head = f.read(4096)
        d = json.loads(head.decode())
        return d if isinstance(d, dict) and STANDIN_KEY in d else None
    except (OSError, ValueError, UnicodeDecodeError):
        return None


def unpack_standin(marker: dict, dest: Path) -> str:
    """Expand an archive marker into the tree the real archive holds (COCO 2017 zips)."""
    dest = Path(dest)
    name = marker["name"]
    stem = name.removesuffix(".zip")
    if stem in COCO_SPLITS:
        write_coco_images(dest, stem)
        return f"{COCO_SPLITS[stem]} synthetic images in {dest / stem}"
    if stem.startswith("annotations_trainval"):
        write_coco_annotations(dest)
        return f"annotations in {dest / 'annotations'}"
    raise SystemExit(f"[b200flow offline] no synthetic stand-in for archive {name}")


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["coco", "redpajama", "standin", "unpack"])
    ap.add_argument("--out", required=True)
    ap.add_argument("--docs", type=int, default=2000)
    ap.add_argument("--url", default="")             # standin: what the download would have fetched
    ap.add_argument("--archive", default="")         # unpack: a marker written by `standin`
    a = ap.parse_args(argv)
    out = Path(a.out)
    if a.kind == "standin":
        print(f"[b200flow offline] wget {a.url} -> {out}: {write_standin(a.url, out, a.docs)}")
        return
    if a.kind == "unpack":
        marker = read_standin(Path(a.archive))
        if marker is None:
            raise SystemExit(f"{a.archive} is not an offline stand-in")
        print(f"[b200flow offline] unzip {a.archive}: {unpack_standin(marker, out)}")
        return
    out.mkdir(parents=True, exist_ok=True)
    if a.kind == "coco":
        write_coco_tree(out)
    else:
        write_standin("book.jsonl", out / "data.jsonl", a.docs)
    print(f"[b200flow] synthetic {a.kind} data written to {out}")


if __name__ == "__main__":
    main()
