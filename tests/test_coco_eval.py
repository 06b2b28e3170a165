"""COCO mAP protocol (b200flow/workloads/maskrcnn/coco_eval.py) on cases whose answer follows from
the definition (pycocotools itself is not in the image)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.workloads.maskrcnn.coco_eval import COCOEvaluator  # noqa: E402


class _Near(float):
    """1.0 up to the `spacing(1)` pycocotools keeps in the precision denominator."""

    def __eq__(self, other):
        return abs(float(self) - float(other)) < 1e-9

    __hash__ = float.__hash__


ONE = _Near(1.0)


def _gt(boxes, labels, **kw):
    return {"boxes": torch.tensor(boxes, dtype=torch.float32).reshape(-1, 4), "labels": torch.tensor(labels), **kw}


def _det(boxes, labels, scores):
    return {"boxes": torch.tensor(boxes, dtype=torch.float32).reshape(-1, 4), "labels": torch.tensor(labels),
            "scores": torch.tensor(scores, dtype=torch.float32)}


def test_perfect_detections_score_one_everywhere():
    ev = COCOEvaluator("bbox")
    ev.add(_det([[10, 10, 110, 110], [200, 50, 260, 90]], [1, 2], [0.9, 0.8]),
           _gt([[10, 10, 110, 110], [200, 50, 260, 90]], [1, 2]))
    s = ev.summarize()
    assert s["AP"] == ONE and s["AP50"] == ONE and s["AP75"] == ONE and s["AR100"] == ONE
    assert s["APl"] == ONE and s["APm"] == ONE and s["APs"] == -1.0      # 100x100 large, 60x40 medium, no small


def test_false_positive_ranked_first_halves_precision():
    ev = COCOEvaluator("bbox")
    ev.add(_det([[300, 300, 400, 400], [10, 10, 110, 110]], [1, 1], [0.9, 0.5]), _gt([[10, 10, 110, 110]], [1]))
    s = ev.summarize()
    assert abs(s["AP"] - 0.5) < 1e-9 and abs(s["AP50"] - 0.5) < 1e-9
    assert s["AR1"] == 0.0 and s["AR10"] == ONE          # the best-scored detection alone finds nothing


def test_localisation_quality_counts_only_below_its_iou():
    # IoU = 62/100 by shrinking the detection's width: matched at thresholds 0.50, 0.55, 0.60 only
    ev = COCOEvaluator("bbox")
    ev.add(_det([[0, 0, 62, 100]], [3], [0.7]), _gt([[0, 0, 100, 100]], [3]))
    s = ev.summarize()
    assert abs(s["AP"] - 0.3) < 1e-9 and s["AP50"] == ONE and s["AP75"] == 0.0


def test_duplicate_detection_is_a_false_positive_and_classes_do_not_mix():
    ev = COCOEvaluator("bbox")
    ev.add(_det([[0, 0, 100, 100], [1, 1, 100, 100], [0, 0, 100, 100]], [1, 1, 2], [0.9, 0.8, 0.95]),
           _gt([[0, 0, 100, 100]], [1]))
    p, _ = ev._accumulate("all", 100)
    # class 1: TP then FP -> precision 1 at every recall point; class 2 has no ground truth -> excluded
    assert ev.summarize()["AP"] == ONE and p.shape[2] == 2 and (p[:, :, 1] == -1).all()


def test_crowd_and_out_of_range_ground_truth_are_ignored_not_penalised():
    ev = COCOEvaluator("bbox")
    gt = _gt([[0, 0, 100, 100], [200, 200, 400, 400]], [1, 1], iscrowd=torch.tensor([0, 1]))
    # two detections inside the crowd region: matched to it, ignored; one true positive
    ev.add(_det([[0, 0, 100, 100], [210, 210, 300, 300], [300, 300, 390, 390]], [1, 1, 1], [0.5, 0.9, 0.8]), gt)
    s = ev.summarize()
    assert s["AP"] == ONE and s["AR100"] == ONE
    # a 20x20 object is "small": its detection must not hurt the "large" number
    ev2 = COCOEvaluator("bbox")
    ev2.add(_det([[0, 0, 20, 20], [50, 50, 250, 250]], [1, 1], [0.9, 0.8]), _gt([[0, 0, 20, 20], [50, 50, 250, 250]], [1, 1]))
    s2 = ev2.summarize()
    assert s2["APs"] == ONE and s2["APl"] == ONE and s2["APm"] == -1.0


def test_segm_uses_mask_iou_and_multi_image_merge_sorts_by_score():
    H = W = 64
    g = torch.zeros(1, H, W, dtype=torch.uint8)
    g[0, 8:40, 8:40] = 1
    d_good = g.clone()
    d_shift = torch.zeros_like(g)
    d_shift[0, 8:40, 24:56] = 1                      # IoU = 1/3 with g: a miss at every threshold
    ev = COCOEvaluator("segm")
    ev.add({"masks": d_shift, "labels": torch.tensor([1]), "scores": torch.tensor([0.9])},
           {"masks": g, "labels": torch.tensor([1])})
    ev.add({"masks": d_good, "labels": torch.tensor([1]), "scores": torch.tensor([0.6])},
           {"masks": g, "labels": torch.tensor([1])})
    s = ev.summarize()
    # ranked list over both images: FP (0.9), TP (0.6); 2 ground truths -> recall tops out at 0.5 with precision 0.5
    want = 0.5 * 51 / 101
    assert abs(s["AP"] - want) < 1e-9 and abs(s["AR100"] - 0.5) < 1e-9
    names = ev.tensorpack_names()
    assert "mAP(segm)/IoU=0.5:0.95" in names and abs(names["mAP(segm)/IoU=0.5"] - want) < 1e-9


def test_merge_of_rank_shards_equals_single_evaluator():
    torch.manual_seed(0)
    imgs = []
    for _ in range(6):
        n = int(torch.randint(1, 5, (1,)))
        xy = torch.rand(n, 2) * 200
        wh = torch.rand(n, 2) * 150 + 10
        gt = {"boxes": torch.cat([xy, xy + wh], 1), "labels": torch.randint(1, 3, (n,))}
        jit = torch.randn(n, 4) * 8
        det = {"boxes": gt["boxes"] + jit, "labels": gt["labels"].clone(), "scores": torch.rand(n)}
        imgs.append((det, gt))
    one, a, b = COCOEvaluator(), COCOEvaluator(), COCOEvaluator()
    for i, (d, g) in enumerate(imgs):
        one.add(d, g)
        (a if i % 2 == 0 else b).add(d, g)
    a.merge(b._records, b.num_images)
    sa, so = a.summarize(), one.summarize()
    assert a.num_images == 6 and all(abs(sa[k] - so[k]) < 1e-12 for k in so)
    assert 0.0 < so["AP"] < 1.0
