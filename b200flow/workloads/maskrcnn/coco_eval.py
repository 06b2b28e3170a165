"""COCO detection metrics (bbox and segm mAP) for the Mask R-CNN workloads.

The reference evaluates every `TRAIN.EVAL_PERIOD` epochs (examples/maskrcnn/train-maskrcnn-aws.yaml:30;
charts/machine-learning/training/maskrcnn/values.yaml) and upstream that means pycocotools'
`COCOeval` (pinned at containers/aws-samples-maskrcnn/Dockerfile:61): AP averaged over IoU
0.50:0.05:0.95, AP50, AP75, AP by object area, AR at 1/10/100 detections — reported by tensorpack
as `mAP(bbox)/IoU=0.5:0.95`, `mAP(segm)/small`, ....  This module restates that protocol:

* per image and category, detections (best score first, at most maxDets) are matched greedily to
  ground truths at each IoU threshold; a ground truth outside the area range (or a crowd region)
  is "ignore": matching it neither counts for nor against, unmatched detections outside the area
  range are ignored too;
* per category the detections of all images are merged by score, precision is made monotone and
  sampled at 101 recall points;
* a summary number is the mean over thresholds, recall points and categories that have ground truth.

IoU matrices are torch expressions on whatever device the detections live on (boxes: broadcast
min/max; masks: one intersection matrix product); matching and accumulation run in NumPy — this is
the evaluation cold path, once per epoch.
"""
from __future__ import annotations

import numpy as np
import torch

IOU_THRS = np.linspace(0.5, 0.95, 10)
REC_THRS = np.linspace(0.0, 1.0, 101)
AREA_RNG = {"all": (0.0, 1e10), "small": (0.0, 32.0 ** 2), "medium": (32.0 ** 2, 96.0 ** 2), "large": (96.0 ** 2, 1e10)}
MAX_DETS = (1, 10, 100)


def box_iou_matrix(d: torch.Tensor, g: torch.Tensor, crowd: np.ndarray) -> np.ndarray:
    """[D, G] IoU of xyxy boxes; for crowd ground truths the union is the detection's own area."""
    if d.numel() == 0 or g.numel() == 0:
        return np.zeros((d.shape[0], g.shape[0]))
    d, g = d.float(), g.float()
    lt = torch.maximum(d[:, None, :2], g[None, :, :2])
    rb = torch.minimum(d[:, None, 2:], g[None, :, 2:])
    inter = (rb - lt).clamp(min=0).prod(-1)
    ad = (d[:, 2:] - d[:, :2]).clamp(min=0).prod(-1)
    ag = (g[:, 2:] - g[:, :2]).clamp(min=0).prod(-1)
    union = ad[:, None] + ag[None, :] - inter
    c = torch.from_numpy(crowd.astype(bool)).to(union.device)
    union = torch.where(c[None, :], ad[:, None].expand_as(union), union)
    return (inter / union.clamp(min=1e-12)).cpu().double().numpy()


def mask_iou_matrix(d: torch.Tensor, g: torch.Tensor, crowd: np.ndarray) -> np.ndarray:
    """[D, G] IoU of binary masks [n, H, W] (intersection as one matrix product)."""
    if d.shape[0] == 0 or g.shape[0] == 0:
        return np.zeros((d.shape[0], g.shape[0]))
    df, gf = d.flatten(1).float(), g.flatten(1).to(d.device).float()
    inter = df @ gf.t()
    ad, ag = df.sum(1), gf.sum(1)
    union = ad[:, None] + ag[None, :] - inter
    c = torch.from_numpy(crowd.astype(bool)).to(union.device)
    union = torch.where(c[None, :], ad[:, None].expand_as(union), union)
    return (inter / union.clamp(min=1e-12)).cpu().double().numpy()


def _match(ious: np.ndarray, gt_ignore: np.ndarray, gt_crowd: np.ndarray):
    """Greedy COCO matching for every IoU threshold.  `ious` rows = detections (best first), columns =
    ground truths ordered non-ignored first.  Returns dt_match [T, D] (gt index or -1) and dt_ignore."""
    T, D, G = len(IOU_THRS), ious.shape[0], ious.shape[1]
    dtm = -np.ones((T, D), dtype=np.int64)
    dtig = np.zeros((T, D), dtype=bool)
    for ti, thr in enumerate(IOU_THRS):
        taken = np.zeros(G, dtype=bool)
        for di in range(D):
            best, m = min(thr, 1 - 1e-10), -1
            for gi in range(G):
                if taken[gi] and not gt_crowd[gi]:
                    continue
                if m > -1 and not gt_ignore[m] and gt_ignore[gi]:
                    break                       # only ignored ground truths follow: keep the real match
                if ious[di, gi] < best:
                    continue
                best, m = ious[di, gi], gi
            if m >= 0:
                dtm[ti, di], dtig[ti, di] = m, gt_ignore[m]
                taken[m] = True
    return dtm, dtig


class COCOEvaluator:
    """Feed images one by one with `add(...)`, then `summarize()`.

    iou_type "bbox": detections / ground truths give `boxes` [n, 4] xyxy; "segm": `masks` [n, H, W]
    binary.  Both take `labels` [n]; detections `scores` [n]; ground truths optionally `iscrowd` [n]
    and `area` [n] (default: box or mask area, as COCO annotations carry it)."""

    def __init__(self, iou_type: str = "bbox"):
        assert iou_type in ("bbox", "segm")
        self.iou_type = iou_type
        self._records: dict = {}          # (cat, area name) -> list of per-image records
        self.num_images = 0

    @staticmethod
    def _areas(obj: dict, key: str) -> np.ndarray:
        if "area" in obj:
            return obj["area"].detach().cpu().double().numpy()
        if key == "bbox":
            b = obj["boxes"].detach().cpu().double()
            return ((b[:, 2] - b[:, 0]).clamp(min=0) * (b[:, 3] - b[:, 1]).clamp(min=0)).numpy()
        return obj["masks"].flatten(1).sum(1).detach().cpu().double().numpy()

    def add(self, det: dict, gt: dict) -> None:
        self.num_images += 1
        key = "boxes" if self.iou_type == "bbox" else "masks"
        dl = det["labels"].detach().cpu().numpy() if det["labels"].numel() else np.zeros(0, dtype=np.int64)
        gl = gt["labels"].detach().cpu().numpy() if gt["labels"].numel() else np.zeros(0, dtype=np.int64)
        ds = det["scores"].detach().cpu().double().numpy() if det["scores"].numel() else np.zeros(0)
        gcrowd_all = (gt["iscrowd"].detach().cpu().numpy().astype(bool) if "iscrowd" in gt
                      else np.zeros(len(gl), dtype=bool))
        darea_all, garea_all = self._areas(det, self.iou_type), self._areas(gt, self.iou_type)
        for cat in np.union1d(dl, gl):
            di = np.nonzero(dl == cat)[0]
            di = di[np.argsort(-ds[di], kind="mergesort")][:MAX_DETS[-1]]
            gi = np.nonzero(gl == cat)[0]
            if self.iou_type == "bbox":
                ious_full = box_iou_matrix(det[key][torch.from_numpy(di)], gt[key][torch.from_numpy(gi)],
                                           gcrowd_all[gi])
            else:
                ious_full = mask_iou_matrix(det[key][torch.from_numpy(di)], gt[key][torch.from_numpy(gi)],
                                            gcrowd_all[gi])
            for aname, (lo, hi) in AREA_RNG.items():
                gig = gcrowd_all[gi] | (garea_all[gi] < lo) | (garea_all[gi] > hi)
                order = np.argsort(gig, kind="mergesort")          # real ground truths first
                dtm, dtig = _match(ious_full[:, order], gig[order], gcrowd_all[gi][order])
                out_of_range = (darea_all[di] < lo) | (darea_all[di] > hi)
                dtig = dtig | ((dtm < 0) & out_of_range[None, :])
                self._records.setdefault((int(cat), aname), []).append(
                    dict(scores=ds[di], matched=dtm >= 0, ignore=dtig, npos=int((~gig).sum())))

    # ------------------------------------------------------------------ accumulate
    def _accumulate(self, aname: str, max_det: int):
        """precision [T, R, K] and recall [T, K]; -1 where a category has no ground truth."""
        cats = sorted({c for (c, a) in self._records if a == aname})
        T, R = len(IOU_THRS), len(REC_THRS)
        precision = -np.ones((T, R, len(cats)))
        recall = -np.ones((T, len(cats)))
        for k, cat in enumerate(cats):
            recs = self._records[(cat, aname)]
            npos = sum(r["npos"] for r in recs)
            if npos == 0:
                continue
            scores = np.concatenate([r["scores"][:max_det] for r in recs])
            order = np.argsort(-scores, kind="mergesort")
            matched = np.concatenate([r["matched"][:, :max_det] for r in recs], 1)[:, order]
            ignore = np.concatenate([r["ignore"][:, :max_det] for r in recs], 1)[:, order]
            tps = np.cumsum(matched & ~ignore, 1).astype(np.float64)
            fps = np.cumsum(~matched & ~ignore, 1).astype(np.float64)
            for t in range(T):
                tp, fp = tps[t], fps[t]
                nd = len(tp)
                rc = tp / npos
                pr = tp / (fp + tp + np.spacing(1))
                recall[t, k] = rc[-1] if nd else 0.0
                for i in range(nd - 1, 0, -1):                     # precision envelope
                    if pr[i] > pr[i - 1]:
                        pr[i - 1] = pr[i]
                inds = np.searchsorted(rc, REC_THRS, side="left")
                q = np.zeros(R)
                ok = inds < nd
                q[ok] = pr[inds[ok]]
                precision[t, :, k] = q
        return precision, recall

    def summarize(self) -> dict:
        """The twelve COCO summary numbers (AP, AP50, AP75, APs, APm, APl, AR1, AR10, AR100, ARs, ARm, ARl)."""
        def mean_valid(x):
            x = x[x > -1]
            return float(x.mean()) if x.size else -1.0

        out = {}
        p_all, r_all = self._accumulate("all", 100)
        out["AP"] = mean_valid(p_all)
        out["AP50"] = mean_valid(p_all[0])
        out["AP75"] = mean_valid(p_all[5])
        for a, n in (("small", "APs"), ("medium", "APm"), ("large", "APl")):
            out[n] = mean_valid(self._accumulate(a, 100)[0])
        for md in MAX_DETS:
            out[f"AR{md}"] = mean_valid(self._accumulate("all", md)[1])
        for a, n in (("small", "ARs"), ("medium", "ARm"), ("large", "ARl")):
            out[n] = mean_valid(self._accumulate(a, 100)[1])
        return out

    def tensorpack_names(self) -> dict:
        """The scalar names tensorpack's `print_coco_metrics` logs (what the jupyter chart's TensorBoard shows)."""
        s, t = self.summarize(), self.iou_type
        return {f"mAP({t})/IoU=0.5:0.95": s["AP"], f"mAP({t})/IoU=0.5": s["AP50"], f"mAP({t})/IoU=0.75": s["AP75"],
                f"mAP({t})/small": s["APs"], f"mAP({t})/medium": s["APm"], f"mAP({t})/large": s["APl"]}

    # --------------------------------------------------------------- multi-rank
    def merge(self, other_records: dict, num_images: int) -> None:
        for k, v in other_records.items():
            self._records.setdefault(k, []).extend(v)
        self.num_images += num_images

    def gather(self, group=None) -> "COCOEvaluator":
        """Every rank evaluated its shard of the validation set: collect the per-image records on all
        ranks (Horovod's eval-result gather)."""
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return self
        parts = [None] * dist.get_world_size(group)
        dist.all_gather_object(parts, (self._records, self.num_images), group=group)
        merged = COCOEvaluator(self.iou_type)
        for recs, n in parts:
            merged.merge(recs, n)
        return merged
