"""Inference entry point with tensorpack's `predict.py` modes:

    python -m b200flow.workloads.maskrcnn.predict --load <logdir | model-N.pt> --evaluate results.json \
        --config DATA.BASEDIR=/fsx/data/coco2017 'DATA.VAL=("coco_val2017")'
    python -m b200flow.workloads.maskrcnn.predict --load <logdir> --predict a.pt b.jpg --out <dir>

`--evaluate` runs the detector over the validation split (sharded over ranks when launched under the launcher's mpirun
mode), writes the detections in COCO results format — `[{image_id, category_id, bbox [x, y, w, h], score, segmentation:
{size, counts}}]` with boxes and masks mapped back to the original image size, masks run-length encoded
(`ops.detection.rle_encode`) — and prints the COCO bbox / segm metrics (`coco_eval.py`).  `--predict` writes one PNG per
image with boxes and pasted masks, like the notebooks of the *-jupyter charts
(containers/tensorpack-maskrcnn/notebooks/mask-rcnn-tensorpack-viz.ipynb:251-269).
"""
from __future__ import annotations

import argparse
import json
import os
from pathlib import Path

import torch
import torch.distributed as dist
import torch.nn.functional as F

from b200flow.models.maskrcnn import MaskRCNN, MaskRCNNConfig, paste_masks
from b200flow.ops import detection as D
from b200flow.workloads.maskrcnn.train import parse_config
from b200flow.workloads.maskrcnn.visualize import draw, latest_checkpoint, write_png


def load_model(load: str, score_thresh: float | None = None):
    src = Path(load)
    if src.is_dir():
        d = src / "train_log" / "maskrcnn" if (src / "train_log" / "maskrcnn").is_dir() else src
        src = latest_checkpoint(d)
        if src is None:
            raise SystemExit(f"no model-*.index under {d}")
    sd = torch.load(src, map_location="cpu", weights_only=False)
    mc = MaskRCNNConfig(**{k: v for k, v in sd.get("model_config", {}).items() if k in MaskRCNNConfig.__dataclass_fields__})
    if score_thresh is not None:
        mc.det_score_thresh = score_thresh
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    model = MaskRCNN(mc)
    model.load_state_dict(sd["model"])
    model.to(dev).eval()
    if dev.type == "cuda":
        model.to(torch.bfloat16).to(memory_format=torch.channels_last)
    return model, mc, dev, src


def detect(model, dev, image: torch.Tensor) -> dict:
    """One image [3, H, W] (padded to a multiple of 32) -> detections with full-size binary masks."""
    h, w = image.shape[-2:]
    H, W = (h + 31) // 32 * 32, (w + 31) // 32 * 32
    x = torch.zeros(1, 3, H, W)
    x[0, :, :h, :w] = image
    x = x.to(dev)
    with torch.no_grad():
        det = model(x.to(torch.bfloat16) if dev.type == "cuda" else x)[0]
    out = {"boxes": det["boxes"].float().cpu(), "scores": det["scores"].float().cpu(), "labels": det["labels"].cpu()}
    if "masks" in det:
        out["masks"] = paste_masks(det, H, W).cpu()[:, :h, :w]
    return out


def evaluate(args, cfgd) -> dict:
    from b200flow.data.coco import COCODetection
    from b200flow.workloads.maskrcnn.coco_eval import COCOEvaluator

    rank = int(os.environ.get("RANK", os.environ.get("OMPI_COMM_WORLD_RANK", 0)))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_SIZE", 1)))
    if world > 1 and not dist.is_initialized():
        from b200flow.parallel.groups import init_distributed

        init_distributed("nccl" if torch.cuda.is_available() else "gloo")
    model, mc, dev, src = load_model(args.load, args.score_thresh)
    split = str(cfgd.get("DATA.VAL", "val2017")).strip("[]()'\", ")
    split = split[len("coco_"):] if split.startswith("coco_") else split
    short, longest = (int(v) for v in args.image_size.split("x"))
    ds = COCODetection(cfgd.get("DATA.BASEDIR", "."), split, train=False, short_side=short, max_side=longest,
                       with_masks=mc.mode_mask)
    evs = {"bbox": COCOEvaluator("bbox")}
    if mc.mode_mask:
        evs["segm"] = COCOEvaluator("segm")
    results = []
    limit = len(ds) if args.max_images is None else min(len(ds), args.max_images)
    for i in range(rank, limit, world):
        s = ds[i]
        det = detect(model, dev, s["image"])
        gt = {k: s[k] for k in ("boxes", "labels", "iscrowd")}
        evs["bbox"].add(det, gt)
        if mc.mode_mask:
            evs["segm"].add({"masks": det.get("masks", torch.zeros(0, *s["image"].shape[-2:], dtype=torch.uint8)),
                             "labels": det["labels"], "scores": det["scores"]}, {**gt, "masks": s["masks"]})
        h0, w0 = s["orig_size"]
        for k in range(det["labels"].numel()):       # COCO results: original image coordinates, category ids
            x1, y1, x2, y2 = (det["boxes"][k] / s["scale"]).tolist()
            rec = {"image_id": s["image_id"], "category_id": ds.label_to_cat.get(int(det["labels"][k]), int(det["labels"][k])),
                   "bbox": [round(x1, 2), round(y1, 2), round(x2 - x1, 2), round(y2 - y1, 2)],
                   "score": round(float(det["scores"][k]), 5)}
            if "masks" in det:
                m = F.interpolate(det["masks"][k][None, None].float(), size=(h0, w0), mode="nearest")[0, 0] > 0.5
                rec["segmentation"] = D.rle_encode(m)
            results.append(rec)
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, results)
        results = [r for p in parts for r in p]
    metrics = {}
    for ev in evs.values():
        metrics.update(ev.gather().tensorpack_names())
    if rank == 0:
        Path(args.evaluate).parent.mkdir(parents=True, exist_ok=True)
        Path(args.evaluate).write_text(json.dumps(results))
        print(f"[b200flow] {len(results)} detections on {limit} images of {split} -> {args.evaluate} (checkpoint {src.name})")
        for k, v in metrics.items():
            print(f"{k}: {v:.4f}")
    if world > 1:
        dist.barrier()
    return metrics


def predict(args) -> None:
    model, mc, dev, src = load_model(args.load, args.score_thresh)
    out = Path(args.out or "predictions")
    out.mkdir(parents=True, exist_ok=True)
    for f in args.predict:
        p = Path(f)
        if p.suffix == ".pt":
            img = torch.load(p).float()
        else:
            import numpy as np
            from PIL import Image

            img = torch.from_numpy(np.asarray(Image.open(p).convert("RGB"), dtype="float32") / 255.0).permute(2, 0, 1)
        det = detect(model, dev, img)
        masks = det.get("masks", torch.zeros(0))
        write_png(out / (p.stem + ".png"), draw(img, det, masks))
        print(f"[b200flow] {p.name}: {det['labels'].numel()} detections -> {out / (p.stem + '.png')}")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--load", required=True)
    ap.add_argument("--evaluate", default=None, help="write COCO-format results of the validation split to this file")
    ap.add_argument("--predict", nargs="*", default=None, help="image files to run on")
    ap.add_argument("--out", default=None)
    ap.add_argument("--config", nargs="*", default=[])
    ap.add_argument("--score-thresh", type=float, default=None)
    ap.add_argument("--image-size", default="800x1333", help="short side x longest side after resizing")
    ap.add_argument("--max-images", type=int, default=None)
    args = ap.parse_args(argv)
    cfgd = parse_config(args.config)
    if args.evaluate:
        evaluate(args, cfgd)
    if args.predict:
        predict(args)
    if not args.evaluate and not args.predict:
        raise SystemExit("one of --evaluate / --predict is required")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
