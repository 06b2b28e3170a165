"""COCO-format detection data for the Mask R-CNN workloads.

What tensorpack's `register_coco` + `get_train_dataflow` do for the reference
(`DATA.BASEDIR=/fsx/data/coco2017`, `DATA.TRAIN=train2017`, `DATA.VAL=val2017`:
charts/machine-learning/training/maskrcnn/templates/maskrcnn.yaml:120-126):

* read `annotations/instances_<split>.json`, group annotations per image, drop images without a usable
  (non-crowd, non-degenerate) box when training;
* resize so that the short side is 800 and the long side at most 1333 (boxes and polygons scale along),
  flip horizontally with probability 1/2 when training;
* turn polygon segmentations into binary masks — K16 (`ops.detection.poly_to_mask`) on a GPU; RLE
  segmentations through `rle_decode`; an annotation without segmentation gets its box as mask;
* pad every image of a batch to one size that is a multiple of 32 (tensorpack's FPN padding; the
  aws-samples fork's `PREPROC.PREDEFINED_PADDING`).

Images are `.pt` tensors (the offline coco-data chart), or `.jpg` / `.png` files when Pillow can read them.
"""
from __future__ import annotations

import json
from pathlib import Path

import torch
import torch.nn.functional as F

from ..ops import detection as D


class COCODetection(torch.utils.data.Dataset):
    def __init__(self, basedir: str | Path, split: str = "train2017", train: bool = True, short_side: int = 800,
                 max_side: int = 1333, with_masks: bool = True, seed: int = 0, mask_device: str = "cpu"):
        self.base, self.split, self.train = Path(basedir), split, train
        self.short, self.max_side, self.with_masks, self.seed = short_side, max_side, with_masks, seed
        self.mask_device = mask_device
        ann = json.loads((self.base / "annotations" / f"instances_{split}.json").read_text())
        cats = sorted(c["id"] for c in ann["categories"])
        self.cat_to_label = {c: i + 1 for i, c in enumerate(cats)}           # contiguous 1..K, 0 = background
        self.label_to_cat = {v: k for k, v in self.cat_to_label.items()}
        self.class_names = ["BG"] + [c["name"] for c in sorted(ann["categories"], key=lambda c: c["id"])]
        per_image: dict = {}
        for a in ann["annotations"]:
            per_image.setdefault(a["image_id"], []).append(a)
        self.images = []
        for im in ann["images"]:
            objs = [a for a in per_image.get(im["id"], []) if a["bbox"][2] > 1 and a["bbox"][3] > 1]
            if train and not any(not a.get("iscrowd", 0) for a in objs):
                continue                                                        # nothing to learn from
            self.images.append((im, objs))

    def __len__(self) -> int:
        return len(self.images)

    def _load_image(self, im: dict) -> torch.Tensor:
        f = self.base / self.split / im["file_name"]
        if f.suffix == ".pt":
            return torch.load(f).float()
        from PIL import Image

        import numpy as np

        arr = np.asarray(Image.open(f).convert("RGB"), dtype="float32") / 255.0
        return torch.from_numpy(arr).permute(2, 0, 1).contiguous()

    def __getitem__(self, i: int) -> dict:
        im, objs = self.images[i]
        img = self._load_image(im)
        h0, w0 = img.shape[-2:]
        scale = min(self.short / min(h0, w0), self.max_side / max(h0, w0))
        h, w = int(round(h0 * scale)), int(round(w0 * scale))
        if (h, w) != (h0, w0):
            img = F.interpolate(img[None], size=(h, w), mode="bilinear", align_corners=False)[0]
        sx, sy = w / w0, h / h0
        g = torch.Generator().manual_seed(self.seed * 1000003 + i)
        flip = self.train and bool(torch.rand(1, generator=g) < 0.5)
        if self.train:
            objs = [a for a in objs if not a.get("iscrowd", 0)]
        boxes = torch.tensor([[a["bbox"][0] * sx, a["bbox"][1] * sy, (a["bbox"][0] + a["bbox"][2]) * sx,
                               (a["bbox"][1] + a["bbox"][3]) * sy] for a in objs], dtype=torch.float32).reshape(-1, 4)
        labels = torch.tensor([self.cat_to_label[a["category_id"]] for a in objs], dtype=torch.int64)
        crowd = torch.tensor([int(a.get("iscrowd", 0)) for a in objs], dtype=torch.int64)
        if flip:
            img = img.flip(-1)
            boxes = torch.stack([w - boxes[:, 2], boxes[:, 1], w - boxes[:, 0], boxes[:, 3]], 1) if len(objs) else boxes
        out = {"image": img, "boxes": boxes, "labels": labels, "iscrowd": crowd, "image_id": im["id"],
               "orig_size": (h0, w0), "scale": scale}
        if self.with_masks:
            out["masks"] = self._masks(objs, boxes, h, w, sx, sy, flip)
        return out

    def _masks(self, objs, boxes, h, w, sx, sy, flip) -> torch.Tensor:
        polys, poly_rows, dense = [], [], {}
        for k, a in enumerate(objs):
            seg = a.get("segmentation")
            if isinstance(seg, list) and seg:
                scaled = []
                for p in seg:
                    pts = torch.tensor(p, dtype=torch.float32).reshape(-1, 2) * torch.tensor([sx, sy])
                    if flip:
                        pts[:, 0] = w - pts[:, 0]
                    scaled.append(pts.reshape(-1).tolist())
                polys.append(scaled)
                poly_rows.append(k)
            elif isinstance(seg, dict) and "counts" in seg:                      # RLE at the original size
                m = D.rle_decode(seg).float()[None, None]
                m = F.interpolate(m, size=(h, w), mode="nearest")[0, 0] > 0.5
                dense[k] = (m.flip(-1) if flip else m).to(torch.uint8)
            else:                                                                 # no segmentation: the box itself
                m = torch.zeros(h, w, dtype=torch.uint8)
                x0, y0, x1, y1 = (int(round(v)) for v in boxes[k].tolist())
                m[max(y0, 0):y1, max(x0, 0):x1] = 1
                dense[k] = m
        masks = torch.zeros(len(objs), h, w, dtype=torch.uint8)
        if polys:
            pm = D.poly_to_mask(polys, h, w, device=self.mask_device).cpu()
            masks[torch.tensor(poly_rows)] = pm
        for k, m in dense.items():
            masks[k] = m
        return masks


def collate_padded(samples: list[dict], multiple: int = 32) -> tuple[torch.Tensor, list[dict]]:
    """Images (and masks) of a batch padded at the bottom / right to one size that is a multiple of
    `multiple`: returns the image batch [n, 3, H, W] and the per-image targets."""
    H = max(s["image"].shape[-2] for s in samples)
    W = max(s["image"].shape[-1] for s in samples)
    H, W = (H + multiple - 1) // multiple * multiple, (W + multiple - 1) // multiple * multiple
    imgs = torch.zeros(len(samples), 3, H, W)
    targets = []
    for i, s in enumerate(samples):
        h, w = s["image"].shape[-2:]
        imgs[i, :, :h, :w] = s["image"]
        t = {k: v for k, v in s.items() if k != "image"}
        if "masks" in s:
            m = torch.zeros(s["masks"].shape[0], H, W, dtype=torch.uint8)
            m[:, :h, :w] = s["masks"]
            t["masks"] = m
        targets.append(t)
    return imgs, targets
