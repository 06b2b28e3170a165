"""Inference + visualisation job of the *-jupyter charts, headless.

The reference's notebooks (containers/tensorpack-maskrcnn/notebooks/
mask-rcnn-tensorpack-viz.ipynb:65,103-105,132,251-252,268-269) glob the newest
`model-*.index` under `$LOGDIR/train_log/maskrcnn`, build the model, run `predict_image` on a
test image and draw boxes + pasted masks with matplotlib.  Here: same checkpoint discovery,
prediction through the b200flow model, masks pasted by the K15 kernel, boxes and masks
rendered into a PNG written by a tiny zlib PNG encoder (no graphics stack), detections
dumped as JSON.  TensorBoard event files written by the trainer sit in the same directory.
"""
from __future__ import annotations

import argparse
import json
import struct
import zlib
from pathlib import Path

import torch

from b200flow.data.synthetic import coco_like
from b200flow.models.maskrcnn import MaskRCNN, MaskRCNNConfig, paste_masks


def write_png(path, rgb: torch.Tensor) -> None:
    """rgb: uint8 [H, W, 3]."""
    h, w, _ = rgb.shape
    raw = b"".join(b"\x00" + rgb[y].contiguous().numpy().tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    Path(path).write_bytes(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                           chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def draw(image: torch.Tensor, det: dict, masks: torch.Tensor) -> torch.Tensor:
    img = (image.float().clamp(0, 1).permute(1, 2, 0) * 255).to(torch.uint8).cpu().clone()
    h, w, _ = img.shape
    palette = torch.tensor([[230, 25, 75], [60, 180, 75], [0, 130, 200], [245, 130, 48], [145, 30, 180],
                            [70, 240, 240], [240, 50, 230], [210, 245, 60]], dtype=torch.float32)
    for i in range(det["boxes"].shape[0]):
        col = palette[i % len(palette)]
        if masks.numel():
            m = masks[i].bool().cpu()
            img[m] = (img[m].float() * 0.5 + col * 0.5).to(torch.uint8)
        x1, y1, x2, y2 = (int(v) for v in det["boxes"][i].round().clamp(min=0).tolist())
        x2, y2 = min(x2, w - 1), min(y2, h - 1)
        img[y1:y2 + 1, [x1, x2]] = col.to(torch.uint8)
        img[[y1, y2], x1:x2 + 1] = col.to(torch.uint8)
    return img


def latest_checkpoint(train_dir: Path) -> Path | None:
    idx = sorted(train_dir.glob("model-*.index"), key=lambda p: int(p.stem.split("-")[1]))
    return (train_dir / idx[-1].read_text().strip()) if idx else None


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--logdir", required=True)
    ap.add_argument("--out", default=None)
    ap.add_argument("--images", type=int, default=2)
    ap.add_argument("--score-thresh", type=float, default=0.05)
    a = ap.parse_args(argv)
    train_dir = Path(a.logdir)
    if (train_dir / "train_log" / "maskrcnn").is_dir():
        train_dir = train_dir / "train_log" / "maskrcnn"
    ck = latest_checkpoint(train_dir)
    if ck is None:
        raise SystemExit(f"no model-*.index under {train_dir}")
    sd = torch.load(ck, map_location="cpu", weights_only=False)
    mc = MaskRCNNConfig(**{k: v for k, v in sd.get("model_config", {}).items() if k in MaskRCNNConfig.__dataclass_fields__})
    mc.det_score_thresh = a.score_thresh
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    model = MaskRCNN(mc)
    model.load_state_dict(sd["model"])
    model.to(dev).eval()
    if dev.type == "cuda":
        model.to(torch.bfloat16).to(memory_format=torch.channels_last)
    out = Path(a.out or (train_dir / "visualization"))
    out.mkdir(parents=True, exist_ok=True)
    H, W = (128, 160) if mc.width < 1.0 else (800, 1344)
    report = []
    for i, s in enumerate(coco_like(a.images, H, W, seed=4242)):
        img = s["image"].to(dev)
        with torch.no_grad():
            det = model((img.to(torch.bfloat16) if dev.type == "cuda" else img)[None])[0]
        masks = paste_masks(det, H, W)
        write_png(out / f"detections_{i}.png", draw(s["image"], det, masks))
        report.append({"image": i, "boxes": det["boxes"].float().cpu().tolist(),
                       "scores": det["scores"].float().cpu().tolist(), "labels": det["labels"].cpu().tolist(),
                       "mask_pixels": [int(m.sum()) for m in masks]})
    (out / "detections.json").write_text(json.dumps({"checkpoint": str(ck), "results": report}, indent=1))
    print(f"[b200flow] wrote {len(report)} visualisations to {out} (checkpoint {ck.name})")


if __name__ == "__main__":
    main()
