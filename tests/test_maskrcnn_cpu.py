"""Mask R-CNN through the launcher on CPU/gloo at test size: the `maskrcnn` chart (mpirun mode, two
ranks, tensorpack command line: charts/machine-learning/training/maskrcnn/templates/maskrcnn.yaml:54-132)
followed by the `maskrcnn-jupyter` chart (charts/machine-learning/testing/maskrcnn-jupyter), which finds
the newest `model-*.index`, predicts and renders boxes + pasted masks headlessly."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
TINY = {"B2F_MRCNN_WIDTH": "0.125", "B2F_MRCNN_IMAGE": "128x128", "CUDA_VISIBLE_DEVICES": ""}


def _install(release, chart, values, root, env):
    cmd = [sys.executable, "-m", "b200flow.launcher", "install", release, chart, "-f", str(values),
           "--root", str(root), "--max-restarts", "0"]
    res = subprocess.run(cmd, cwd=REPO, env=dict(os.environ, **env), capture_output=True, text=True, timeout=900)
    return res.returncode, json.loads(res.stdout.strip().splitlines()[-1]), res.stdout[-2000:] + res.stderr[-2000:]


def test_maskrcnn_chart_then_jupyter_chart(tmp_root):
    v = tmp_root / "train.yaml"
    v.write_text("maskrcnn:\n  gpus: 2\n  gpu_nodes: 1\n  gpus_per_node: 2\n  steps_per_epoch: 2\n"
                 "  lr_schedule: \"[4,6,8]\"\n")
    rc, st, out = _install("mrcnn", "maskrcnn", v, tmp_root, dict(TINY, B2F_MRCNN_EPOCHS="2"))
    assert rc == 0 and st["phase"] == "Succeeded" and len(st["pods"]) == 2, out
    logdirs = list((tmp_root / "fsx/logs").iterdir())
    assert len(logdirs) == 1
    ck = logdirs[0] / "train_log/maskrcnn"
    names = {p.name for p in ck.iterdir()}
    # tensorpack ModelSaver layout + TensorBoard events + config dump (TRAIN.CHECKPOINT_PERIOD=2 from the chart)
    assert {"checkpoint", "model-4.index", "model-4.pt", "config.json", "metrics.jsonl"} <= names
    assert any(n.startswith("events.out.tfevents") for n in names)
    # TRAIN.EVAL_PERIOD=1: COCO bbox and segm mAP (tensorpack's scalar names) after every epoch, gathered over ranks
    metrics = [json.loads(l) for l in (ck / "metrics.jsonl").read_text().splitlines()]
    evals = [m for m in metrics if "mAP(bbox)/IoU=0.5:0.95" in m]
    assert len(evals) == 2 and all("mAP(segm)/IoU=0.5" in m and -1.0 <= m["mAP(bbox)/IoU=0.5"] <= 1.0 for m in evals)
    cfg = json.loads((ck / "config.json").read_text())
    assert cfg["world"] == 2 and cfg["config"]["MODE_MASK"] is True and cfg["config"]["TRAIN.CHECKPOINT_PERIOD"] == 2

    j = tmp_root / "viz.yaml"
    j.write_text(f"global:\n  log_dir: logs/{logdirs[0].name}\n")
    rc, st, out = _install("viz", "maskrcnn-jupyter", j, tmp_root, TINY)
    assert rc == 0 and st["phase"] == "Succeeded", out
    viz = logdirs[0] / "visualization"
    pngs = sorted(viz.glob("detections_*.png"))
    assert pngs and pngs[0].read_bytes()[:8] == b"\x89PNG\r\n\x1a\n"
    dets = json.loads((viz / "detections.json").read_text())
    assert isinstance(dets, (list, dict)) and len(dets) > 0


def test_tensorpack_npz_backbone_loads_into_resnet50(tmp_path):
    """BACKBONE.WEIGHTS: a tensorpack-named ResNet-50 npz (HWIO convolutions, bn gamma/beta/EMA statistics)
    fills every convolution and norm tensor of the body; the offline placeholder is recognised."""
    import numpy as np
    import torch
    import torchvision
    from torchvision.ops import FrozenBatchNorm2d

    from b200flow.models.maskrcnn import load_tensorpack_backbone, tensorpack_to_torchvision_names

    torch.manual_seed(0)
    src = torchvision.models.resnet50(weights=None, norm_layer=FrozenBatchNorm2d)
    for k, v in src.state_dict().items():
        if "bn" in k or "downsample.1" in k:
            v.copy_(torch.rand_like(v) + 0.5)
    table = tensorpack_to_torchvision_names()
    sd = src.state_dict()
    body_keys = {k for k in sd if not k.startswith("fc.")}
    assert {tv for tv, _ in table.values()} == body_keys           # the table covers the whole body, nothing else
    arrays = {tp: (sd[tv].permute(2, 3, 1, 0) if kind == "conv" else sd[tv]).numpy() for tp, (tv, kind) in table.items()}
    arrays["linear/W"] = np.zeros((2048, 1000), dtype=np.float32)   # classifier head: present upstream, unused here
    np.savez(tmp_path / "ImageNet-R50-AlignPadding.npz", **arrays)
    dst = torchvision.models.resnet50(weights=None, norm_layer=FrozenBatchNorm2d)
    rep = load_tensorpack_backbone(dst, tmp_path / "ImageNet-R50-AlignPadding.npz")
    assert rep == {"loaded": len(table), "missing": [], "unused": ["linear/W"], "placeholder": False}
    for k in body_keys:
        assert torch.equal(dst.state_dict()[k], sd[k]), k
    np.savez(tmp_path / "placeholder.npz", synthetic=np.zeros(1))
    assert load_tensorpack_backbone(dst, tmp_path / "placeholder.npz")["placeholder"] is True


def test_training_reads_the_coco_tree_written_by_the_coco_data_chart(tmp_root):
    """coco-data chart (offline: synthetic COCO-format tree with polygon segmentations) -> maskrcnn chart with
    its default DATA.BASEDIR: annotations are parsed, polygons rasterised, images resized / flipped / padded,
    and the validation split is scored with the COCO protocol."""
    import torch

    from b200flow.data.coco import COCODetection, collate_padded

    v = tmp_root / "data.yaml"
    v.write_text("{}\n")
    rc, st, out = _install("d-coco", "coco-data", v, tmp_root, TINY)
    assert rc == 0, out
    base = tmp_root / "fsx/data/coco2017"
    ds = COCODetection(base, "train2017", train=True, short_side=96, max_side=128, seed=3)
    assert len(ds) == 16
    s = ds[0]
    h, w = s["image"].shape[-2:]
    assert min(h, w) <= 96 and max(h, w) <= 128 and s["masks"].shape == (len(s["labels"]), h, w)
    for b, m in zip(s["boxes"], s["masks"]):                      # a rasterised ellipse fills ~pi/4 of its box
        x0, y0, x1, y1 = b.tolist()
        frac = float(m.sum()) / max(1.0, (x1 - x0) * (y1 - y0))
        assert 0.55 < frac < 1.0 and float(m[:, : max(0, int(x0) - 1)].sum()) == 0
    flipped = [i for i in range(len(ds)) if not torch.equal(ds[i]["boxes"], COCODetection(
        base, "train2017", train=False, short_side=96, max_side=128)[i]["boxes"])]
    assert 0 < len(flipped) < len(ds)                              # the training view flips about half the images
    imgs, tg = collate_padded([ds[0], ds[1]])
    assert imgs.shape[-1] % 32 == 0 and imgs.shape[-2] % 32 == 0 and tg[0]["masks"].shape[-2:] == imgs.shape[-2:]

    t = tmp_root / "train.yaml"
    t.write_text("maskrcnn:\n  gpus: 2\n  gpu_nodes: 1\n  gpus_per_node: 2\n  steps_per_epoch: 2\n"
                 "  lr_schedule: \"[4,6,8]\"\n")
    rc, st, out = _install("mrcnn-coco", "maskrcnn", t, tmp_root, dict(TINY, B2F_MRCNN_EPOCHS="2", B2F_DATA_WORKERS="2"))
    logs = "\n".join(p.read_text(errors="replace") for p in
                     (tmp_root / ".b200flow/releases/mrcnn-coco/pods").glob("*/stdout.log"))
    assert rc == 0 and st["phase"] == "Succeeded", logs[-3000:]
    assert "16 training images, 4 validation images" in logs and "mAP(segm)/IoU=0.5:0.95" in logs
    assert "BACKBONE.WEIGHTS" not in logs or "placeholder" in logs
    # tensorpack's predict.py modes on the checkpoint just written: COCO results file + metrics, and a rendered image
    import subprocess

    logdir = next((tmp_root / "fsx/logs").iterdir())
    env = dict(os.environ, **TINY, PYTHONPATH=str(REPO))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    res = subprocess.run([sys.executable, "-m", "b200flow.workloads.maskrcnn.predict", "--load", str(logdir),
                          "--evaluate", str(tmp_root / "out" / "results.json"), "--image-size", "96x128",
                          "--score-thresh", "0.0", "--config", f"DATA.BASEDIR={base}", 'DATA.VAL=("coco_val2017")',
                          "--predict", str(base / "val2017" / "000000000000.pt"), "--out", str(tmp_root / "out")],
                         cwd=REPO, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, (res.stdout + res.stderr)[-3000:]
    assert "mAP(bbox)/IoU=0.5:0.95" in res.stdout and "mAP(segm)/IoU=0.5" in res.stdout
    results = json.loads((tmp_root / "out" / "results.json").read_text())
    assert results and {"image_id", "category_id", "bbox", "score", "segmentation"} <= set(results[0])
    r0 = results[0]
    assert r0["segmentation"]["size"] == [128, 160] and sum(r0["segmentation"]["counts"]) == 128 * 160
    assert 1 <= r0["category_id"] <= 80 and len(r0["bbox"]) == 4 and r0["bbox"][2] > 0
    assert (tmp_root / "out" / "000000000000.png").read_bytes()[:8] == b"\x89PNG\r\n\x1a\n"


def test_stock_library_twin_of_maskrcnn_builds_and_steps_on_cpu():
    """`bench.py --impl twin --workload maskrcnn`: torchvision's Mask R-CNN on the same synthetic stream.  The
    model / batch / loss pieces are checked here at a small image size (the timing harness needs a GPU)."""
    import torch

    from b200flow.workloads.twin import maskrcnn_twin

    torch.manual_seed(0)
    model, make, loss_fn = maskrcnn_twin(torch.device("cpu"), torch.float32, batch=2, H=128, W=160, seed=5)
    b = make()
    assert b["images"].shape == (2, 3, 128, 160) and b["masks0"].shape[1:] == (128, 160)
    model.train()
    loss = loss_fn(model, b)
    loss.backward()
    assert torch.isfinite(loss) and any(p.grad is not None and p.grad.abs().sum() > 0 for p in model.parameters())


def test_faster_rcnn_mode_without_mask_head(tmp_path):
    """`MODE_MASK=False` (tensorpack's Faster R-CNN configuration of the same script): no mask head, no mask loss,
    bbox metrics only."""
    import subprocess

    env = dict(os.environ, **TINY, B2F_MRCNN_EPOCHS="1", B2F_MRCNN_STEPS="2", PYTHONPATH=str(REPO))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    res = subprocess.run([sys.executable, "-m", "b200flow.workloads.maskrcnn.train", "--logdir", str(tmp_path / "log"),
                          "--config", "MODE_MASK=False", "MODE_FPN=True", "TRAIN.EVAL_PERIOD=1"],
                         cwd=REPO, env=env, capture_output=True, text=True, timeout=600)
    out = res.stdout + res.stderr
    assert res.returncode == 0, out[-2000:]
    assert "mAP(bbox)/IoU=0.5:0.95" in out and "mAP(segm)" not in out and "mask" not in out.split("[epoch 1]")[1].split("|")[1]
    again = subprocess.run([sys.executable, "-m", "b200flow.workloads.maskrcnn.train", "--logdir", str(tmp_path / "log2"),
                            "--load", str(tmp_path / "log"), "--config", "MODE_MASK=False", "TRAIN.LR_SCHEDULE=1x"],
                           cwd=REPO, env=env, capture_output=True, text=True, timeout=600)
    assert again.returncode == 0 and "initial weights from" in again.stdout and "resumed from" not in again.stdout, \
        (again.stdout + again.stderr)[-2000:]
    refused = subprocess.run([sys.executable, "-m", "b200flow.workloads.maskrcnn.train", "--logdir", str(tmp_path / "c4"),
                              "--config", "MODE_FPN=False"], cwd=REPO, env=env, capture_output=True, text=True, timeout=600)
    assert refused.returncode != 0 and "MODE_FPN=False" in refused.stderr


@pytest.mark.parametrize("values,expect", [
    ("maskrcnn/train-maskrcnn-tensorpack.yaml", {"batch": 1}),       # W6: tensorpack, 1 image per GPU
    ("maskrcnn/train-maskrcnn-aws.yaml", {"batch": 4}),              # W7: aws-samples fork, TRAIN.BATCH_SIZE_PER_GPU=4
])
def test_reference_mpijob_values_files_run_unchanged(tmp_root, values, expect):
    """The two Mask R-CNN example values files through the `mpijob-horovod-tensorflow-gpu` chart in the launcher's
    mpirun mode (2 ranks), including their `cd /mask-rcnn-tensorflow` line that fails harmlessly as it does upstream."""
    ex = Path("/root/reference/examples")
    if not ex.exists():
        pytest.skip("reference tree not mounted")
    cmd = [sys.executable, "-m", "b200flow.launcher", "install", "w", "mpijob-horovod-tensorflow-gpu", "-f", str(ex / values),
           "--root", str(tmp_root), "--max-restarts", "0", "--set", "resources.gpu_nodes=1", "--set", "resources.gpus_per_node=2"]
    env = dict(os.environ, **TINY, B2F_MRCNN_EPOCHS="1", B2F_MRCNN_STEPS="2")
    res = subprocess.run(cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=900)
    st = json.loads(res.stdout.strip().splitlines()[-1])
    logs = "\n".join(p.read_text(errors="replace") for p in (tmp_root / ".b200flow/releases/w/pods").glob("*/stdout.log"))
    assert res.returncode == 0 and st["phase"] == "Succeeded" and len(st["pods"]) == 2, logs[-3000:]
    assert "[epoch 1] step 2" in logs and "mAP(bbox)/IoU=0.5:0.95" in logs, logs[-3000:]
    import re

    thr = float(re.search(r"throughput=([0-9.]+) images/s", logs).group(1))
    assert thr > 0
    cfg = json.loads(next((tmp_root / "efs/home/w/logs").rglob("config.json")).read_text()) if list(
        (tmp_root / "efs/home/w/logs").rglob("config.json")) else None
    if cfg is not None:
        assert int(cfg["config"].get("TRAIN.BATCH_SIZE_PER_GPU", 1)) == expect["batch"] and cfg["world"] == 2


def test_fixed_shape_losses_stay_finite_with_unused_slots():
    """Unused sampler slots may hold degenerate boxes (zero width -> inf / nan box targets); they must be selected out
    of the losses, not multiplied by zero, and a randomly initialised backbone must start near the chance losses."""
    import torch

    from b200flow.data.synthetic import coco_like
    from b200flow.models import maskrcnn as M
    from b200flow.optim import FlatMomentumSGD
    from b200flow.utils.flat import FlatParams

    assert float(M._masked(torch.tensor([float("nan"), float("inf"), 2.0]), torch.tensor([False, False, True])).sum()) == 2.0
    a = torch.tensor([[5.0, 5.0, 5.0, 5.0]])                       # a zero-size proposal
    bad = M.encode_boxes(torch.tensor([[0.0, 0.0, 4.0, 4.0]]), a, (1.0, 1.0, 1.0, 1.0))
    assert not torch.isfinite(bad).all()
    assert torch.equal(M._masked(bad, torch.tensor([False])), torch.zeros(1, 4))

    torch.manual_seed(0)
    model = M.MaskRCNN(M.MaskRCNNConfig(width=0.25))
    model.static_shapes = True
    flat = FlatParams(model)
    opt = FlatMomentumSGD(flat, lr=0.005, momentum=0.9, weight_decay=1e-4)
    model.train()
    data = coco_like(10 ** 9, 128, 160, seed=3, device="cpu")
    totals = []
    for _ in range(4):
        s = [next(data) for _ in range(2)]
        im = torch.stack([x["image"] for x in s])
        tg = [{k: x[k] for k in ("boxes", "labels", "masks")} for x in s]
        opt.zero_grad()
        losses = model(im, tg)
        losses["total"].backward()
        flat.collect_autograd_grads()
        opt.step()
        assert all(torch.isfinite(v) for v in losses.values()), losses
        totals.append(float(losses["total"].detach()))
    assert totals[0] < 12.0, totals          # ln(81) + ln 2 + ln 2 + box terms; the old initialisation started at 20-40


def test_batched_fixed_shape_sampler_and_padded_targets():
    """The fixed-shape sampler over [N, n] masks samples every row independently, inside the quotas; padded targets
    carry zero boxes that match nothing; an image without objects contributes negatives only."""
    import torch

    from b200flow.data.synthetic import coco_like
    from b200flow.models import maskrcnn as M

    torch.manual_seed(1)
    pos = torch.zeros(3, 50, dtype=torch.bool)
    pos[0, :5] = True            # fewer positives than the quota
    pos[1, :40] = True           # more than the quota of 8
    neg = ~pos
    neg[2] = False               # nothing at all in row 2
    pi, pok, ni, nok = M._sample_static(pos, neg, 16, 0.5)
    assert pi.shape == (3, 8) and ni.shape == (3, 16)
    assert pok.sum(1).tolist() == [5, 8, 0] and nok.sum(1).tolist() == [11, 8, 0]
    for r in range(3):
        assert pos[r, pi[r][pok[r]]].all() and neg[r, ni[r][nok[r]]].all()
        assert len(set(pi[r][pok[r]].tolist())) == int(pok[r].sum())          # no duplicates

    data = coco_like(3, 96, 128, seed=5, device="cpu")
    tg = [{k: s[k] for k in ("boxes", "labels", "masks")} for s in data]
    tg[1] = {"boxes": tg[1]["boxes"][:0], "labels": tg[1]["labels"][:0], "masks": tg[1]["masks"][:0]}
    b, l, v = M.pad_targets(tg)
    assert b.shape[0] == 3 and v[1].sum() == 0 and v[0].sum() == tg[0]["boxes"].shape[0]
    assert torch.equal(b[0][v[0]], tg[0]["boxes"].float()) and (b[~v] == 0).all()

    # fixed-shape and variable-length formulations estimate the same losses: at initialisation they agree closely
    torch.manual_seed(0)
    model = M.MaskRCNN(M.MaskRCNNConfig(width=0.25))
    model.train()
    ims = torch.stack([s["image"] for s in coco_like(3, 96, 128, seed=5, device="cpu")])
    model.static_shapes = True
    ls = model(ims, tg)
    model.static_shapes = False
    ld = model(ims, tg)
    assert all(torch.isfinite(x) for x in ls.values())
    assert abs(float(ls["rpn_cls"]) - float(ld["rpn_cls"])) < 0.05
    assert abs(float(ls["frcnn_cls"]) - float(ld["frcnn_cls"])) < 0.3
    assert abs(float(ls["mrcnn"]) - float(ld["mrcnn"])) < 0.2
