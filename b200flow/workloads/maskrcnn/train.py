"""Mask R-CNN trainer with the command line of the reference's training scripts:

    python3 train.py --logdir <dir> [--images_per_epoch N] --config KEY=VALUE ...

(charts/machine-learning/training/maskrcnn/templates/maskrcnn.yaml:114-132 for tensorpack's
`examples/FasterRCNN/train.py`; maskrcnn-optimized/templates/maskrcnn.yaml:117-136 for
aws-samples' `MaskRCNN/train.py`; launched by mpirun with TRAINER=horovod).  Recognised
config keys: MODE_MASK, MODE_FPN, DATA.BASEDIR, DATA.TRAIN, DATA.VAL, TRAIN.EVAL_PERIOD,
TRAIN.STEPS_PER_EPOCH, TRAIN.LR_SCHEDULE, TRAIN.LR_EPOCH_SCHEDULE, TRAIN.BASE_LR,
TRAIN.BATCH_SIZE_PER_GPU, TRAIN.CHECKPOINT_PERIOD, TRAIN.GRADIENT_CLIP, TRAIN.WARMUP,
BACKBONE.WEIGHTS, BACKBONE.NORM, PREPROC.PREDEFINED_PADDING, TRAINER (others are stored
and ignored).  Ranks come from the launcher's mpirun mode (OMPI_COMM_WORLD_* / RANK).

Callbacks mirror tensorpack's: periodic model saver (`model-<step>.pt` + `model-<step>.index`
marker + `checkpoint` file under `<logdir>`, auto-resume from the newest), LR schedule with
linear warm-up, throughput / memory / time-left trackers, evaluation every EVAL_PERIOD epochs (COCO bbox / segm mAP on the
synthetic validation split, coco_eval.py), TensorBoard scalars.  COCO-shaped synthetic images.

Test-size knobs: B2F_MRCNN_WIDTH, B2F_MRCNN_IMAGE (HxW), B2F_MRCNN_EPOCHS, B2F_MRCNN_STEPS.
"""
from __future__ import annotations

import argparse
import ast
import json
import os
import time
from pathlib import Path

import torch
import torch.distributed as dist

from b200flow.data.synthetic import coco_like
from b200flow.models.maskrcnn import MaskRCNN, MaskRCNNConfig
from b200flow.obs.tracking import Tracker, maybe_inject_fault
from b200flow.optim import FlatMomentumSGD
from b200flow.parallel.groups import init_distributed


def parse_config(items):
    cfg = {}
    for it in items or []:
        k, _, v = it.partition("=")
        try:
            cfg[k] = ast.literal_eval(v)
        except (ValueError, SyntaxError):
            cfg[k] = v
    return cfg


def lr_at(step, base_lr, world, batch, warmup, boundaries):
    """tensorpack scales BASE_LR by total batch / 8 and decays 10x at each boundary (given for
    a total batch of 8); linear warm-up from lr/3."""
    lr = base_lr * (world * batch) / 8.0
    scale = 8.0 / (world * batch)
    if step < warmup:
        return lr * (1 / 3 + (2 / 3) * step / max(1, warmup))
    for i, b in enumerate(boundaries):
        if step < b * scale:
            return lr * (0.1 ** i)
    return lr * (0.1 ** len(boundaries))


def evaluate_coco(model, val, dev, use_cuda, mode_mask):
    """COCO protocol on this rank's validation shard, merged over ranks: tensorpack's
    `mAP(bbox)/...` and `mAP(segm)/...` scalars (coco_eval.py)."""
    from b200flow.models.maskrcnn import paste_masks
    from b200flow.workloads.maskrcnn.coco_eval import COCOEvaluator

    evs = {"bbox": COCOEvaluator("bbox")}
    if mode_mask:
        evs["segm"] = COCOEvaluator("segm")
    with torch.no_grad():
        vi = torch.stack([s["image"] for s in val]).to(dev)
        dets = model(vi.to(torch.bfloat16) if use_cuda else vi)
    H, W = vi.shape[-2:]
    for d, g in zip(dets, val):
        d = {k: v for k, v in d.items() if k != "fastrcnn_all_scores"}
        evs["bbox"].add({"boxes": d["boxes"].float().cpu(), "labels": d["labels"].cpu(), "scores": d["scores"].float().cpu()},
                        {"boxes": g["boxes"].float().cpu(), "labels": g["labels"].cpu(),
                         **({"iscrowd": g["iscrowd"].cpu()} if "iscrowd" in g else {})})
        if mode_mask:
            pm = paste_masks(d, H, W).cpu() if d["labels"].numel() else torch.zeros(0, H, W, dtype=torch.uint8)
            evs["segm"].add({"masks": pm, "labels": d["labels"].cpu(), "scores": d["scores"].float().cpu()},
                            {"masks": g["masks"].cpu(), "labels": g["labels"].cpu(),
                             **({"iscrowd": g["iscrowd"].cpu()} if "iscrowd" in g else {})})
    out = {}
    for ev in evs.values():
        out.update(ev.gather().tensorpack_names())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--logdir", required=True)
    ap.add_argument("--images_per_epoch", type=int, default=None)
    ap.add_argument("--config", nargs="*", default=[])
    ap.add_argument("--load", default=None)
    args, unknown = ap.parse_known_args()
    cfgd = parse_config(args.config)
    use_cuda = torch.cuda.is_available()
    rank, local_rank, world = init_distributed("nccl" if use_cuda else "gloo") if "RANK" in os.environ else (0, 0, 1)
    dev = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    batch = int(cfgd.get("TRAIN.BATCH_SIZE_PER_GPU", 1))
    ipe = args.images_per_epoch
    steps_per_epoch = int(os.environ.get("B2F_MRCNN_STEPS", 0)) or (
        ipe // (world * batch) if ipe else int(cfgd.get("TRAIN.STEPS_PER_EPOCH", 7500)))
    if "TRAIN.LR_EPOCH_SCHEDULE" in cfgd:      # [(16, 0.1), (20, 0.01), (24, None)] in epochs
        sched = cfgd["TRAIN.LR_EPOCH_SCHEDULE"]
        total_epochs = sched[-1][0]
        boundaries = [e * 120000 // 8 for e, _ in sched[:-1]]
    else:
        sched = cfgd.get("TRAIN.LR_SCHEDULE", [240000, 320000, 360000])
        if isinstance(sched, str) and sched.endswith("x"):      # tensorpack's "1x" / "2x" ... = multiples of 90k steps
            mult = float(sched[:-1])
            sched = [int(120000 * mult), int(160000 * mult), int(180000 * mult)]
        boundaries = list(sched)
        total_epochs = max(1, int(boundaries[-1] * 8 / (world * batch)) // max(1, steps_per_epoch))
        boundaries = boundaries[:-1]
    total_epochs = int(os.environ.get("B2F_MRCNN_EPOCHS", total_epochs))
    H, W = (int(v) for v in os.environ.get("B2F_MRCNN_IMAGE", "800x1344").split("x"))
    width = float(os.environ.get("B2F_MRCNN_WIDTH", 1.0))
    mcfg = MaskRCNNConfig(mode_mask=bool(cfgd.get("MODE_MASK", True)), mode_fpn=bool(cfgd.get("MODE_FPN", True)),
                          width=width)
    if width < 1.0:
        mcfg.rpn_pre_nms_train = mcfg.rpn_post_nms_train = 300
        mcfg.rpn_pre_nms_test = mcfg.rpn_post_nms_test = 200
        mcfg.roi_batch = 128
    torch.manual_seed(1234)
    model = MaskRCNN(mcfg).to(dev)
    wpath = str(cfgd.get("BACKBONE.WEIGHTS", "") or "")
    if wpath and os.path.exists(wpath) and width >= 1.0:       # ImageNet ResNet-50 in tensorpack's npz naming
        from b200flow.models.maskrcnn import load_tensorpack_backbone

        rep = load_tensorpack_backbone(model.backbone.body, wpath)
        if rank == 0:
            what = "offline placeholder, backbone stays randomly initialised" if rep["placeholder"] else \
                f"{rep['loaded']} tensors loaded, {len(rep['missing'])} missing, {len(rep['unused'])} unused"
            print(f"[b200flow] BACKBONE.WEIGHTS={wpath}: {what}", flush=True)
    if use_cuda:
        model = model.to(torch.bfloat16).to(memory_format=torch.channels_last)
    ddp, tuner = None, None
    if world > 1:      # TRAINER=horovod: fusion-buffer size / autotune / timeline come from the HOROVOD_* variables
        from b200flow.engines.horovod_api import make_ddp

        ddp, tuner = make_ddp(model)
    if ddp is None:
        from b200flow.utils.flat import FlatParams

        flat = FlatParams(model)
    else:
        flat = ddp.flat
    opt = FlatMomentumSGD(flat, lr=float(cfgd.get("TRAIN.BASE_LR", 0.01)), momentum=0.9, weight_decay=1e-4)
    clip = float(cfgd.get("TRAIN.GRADIENT_CLIP", 0) or 0)
    logdir = Path(args.logdir)
    if rank == 0:
        logdir.mkdir(parents=True, exist_ok=True)
        (logdir / "config.json").write_text(json.dumps({"config": cfgd, "world": world}, default=str, indent=1))
    tracker = Tracker(logdir, ".") if rank == 0 else None
    # ---- auto-resume from <logdir>/checkpoint (tensorpack behaviour)
    step = 0
    ckfile = logdir / "checkpoint"
    if ckfile.exists():
        name = json.loads(ckfile.read_text())["model_checkpoint_path"]
        sd = torch.load(logdir / (name + ".pt"), map_location="cpu", weights_only=False)
        model.load_state_dict(sd["model"])
        opt.load_state_dict({k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in sd["optimizer"].items()})
        step = sd["global_step"]
        if rank == 0:
            print(f"[b200flow] resumed from {name} (step {step})", flush=True)
    elif args.load:          # --load <model-N.pt | logdir>: initial weights only (tensorpack's `--load`), step stays 0
        src = Path(args.load)
        if src.is_dir():
            src = src / (json.loads((src / "checkpoint").read_text())["model_checkpoint_path"] + ".pt")
        sd = torch.load(src, map_location="cpu", weights_only=False)
        model.load_state_dict(sd["model"] if "model" in sd else sd)
        opt.master.copy_(flat.params.float())          # the fp32 master copy follows the loaded weights
        if rank == 0:
            print(f"[b200flow] initial weights from {src}", flush=True)
    warmup = int(cfgd.get("TRAIN.WARMUP", min(1000, steps_per_epoch)))
    ckpt_period = int(cfgd.get("TRAIN.CHECKPOINT_PERIOD", 1))
    eval_period = int(cfgd.get("TRAIN.EVAL_PERIOD", 1))
    data = coco_like(10 ** 9, H, W, seed=1000 + rank, device="cpu")
    # DATA.BASEDIR with COCO annotations (the coco-data chart's output): train on it, evaluate on DATA.VAL
    train_set = val_set = None
    basedir = str(cfgd.get("DATA.BASEDIR", "") or "")
    def split_name(v, default):        # tensorpack registers COCO splits as "coco_train2017": ["coco_train2017"] / ("coco_val2017")
        v = str(v if v else default).strip("[]()'\", ")
        return v[len("coco_"):] if v.startswith("coco_") else v

    split_t, split_v = split_name(cfgd.get("DATA.TRAIN"), "train2017"), split_name(cfgd.get("DATA.VAL"), "val2017")
    if basedir and os.path.exists(os.path.join(basedir, "annotations", f"instances_{split_t}.json")):
        from b200flow.data.coco import COCODetection

        kw = dict(short_side=min(H, W), max_side=max(H, W), with_masks=mcfg.mode_mask)
        train_set = COCODetection(basedir, split_t, train=True, seed=1000, **kw)
        if os.path.exists(os.path.join(basedir, "annotations", f"instances_{split_v}.json")):
            val_set = COCODetection(basedir, split_v, train=False, **kw)
        if rank == 0:
            print(f"[b200flow] DATA.BASEDIR={basedir}: {len(train_set)} training images"
                  + (f", {len(val_set)} validation images" if val_set is not None else ""), flush=True)
    start_epoch = step // max(1, steps_per_epoch)
    graph_tried = False
    for epoch in range(start_epoch + 1, total_epochs + 1):
        model.train()
        t0, seen, agg = time.time(), 0, {}
        if train_set is not None:      # this rank's images of the epoch, produced ahead by worker processes
            from b200flow.data.prefetch import PrefetchLoader

            order = [((step + k) * world + rank) * batch + j for k in range(steps_per_epoch) for j in range(batch)]
            feed = iter(PrefetchLoader(train_set, [i % len(train_set) for i in order],
                                       workers=int(os.environ.get("B2F_DATA_WORKERS", 4 if use_cuda else 0))))
        for _ in range(steps_per_epoch):
            maybe_inject_fault(step)
            if train_set is not None:
                from b200flow.data.coco import collate_padded

                imgs, samples = collate_padded([next(feed) for _ in range(batch)])
                imgs = imgs.to(dev)
            else:
                samples = [next(data) for _ in range(batch)]
                imgs = torch.stack([s["image"] for s in samples]).to(dev)
            if use_cuda:
                imgs = imgs.to(torch.bfloat16)
            tg = [{k: s[k].to(dev) for k in ("boxes", "labels", "masks")} for s in samples]
            opt.param_groups[0]["lr"] = lr_at(step, float(cfgd.get("TRAIN.BASE_LR", 0.01)), world, batch,
                                              warmup, boundaries)
            if use_cuda and not graph_tried:
                # backbone + FPN + RPN head as CUDA graphs once the weights are final and the image shape is known
                graph_tried = True
                if model.static_shapes and os.environ.get("B2F_MASKRCNN_GRAPH", "1") == "1":
                    model.enable_trunk_graph(imgs)
            opt.zero_grad()
            losses = model(imgs, tg)
            losses["total"].backward()
            flat.collect_autograd_grads()
            if ddp is not None:
                ddp.finish_reduction()
            if clip > 0:
                gn = flat.grads.norm()
                opt.grad_scale = (clip / (gn + 1e-6)).clamp(max=1.0).reshape(1)
            opt.step()
            if tuner is not None:
                tuner.step_done()
            step += 1
            seen += batch * world
            for k, v in losses.items():
                agg[k] = agg.get(k, 0.0) + float(v.detach())
        dt = time.time() - t0
        if rank == 0:
            msg = {k: v / steps_per_epoch for k, v in agg.items()}
            print(f"[epoch {epoch}] step {step} | " + " ".join(f"{k}={v:.4f}" for k, v in msg.items()) +
                  f" | lr={opt.param_groups[0]['lr']:.5f} | throughput={seen / dt:.2f} images/s", flush=True)
            # tensorpack's GPUMemoryTracker / HostMemoryTracker / EstimatedTimeLeft callbacks
            import psutil

            gpu_gb = torch.cuda.max_memory_allocated() / 2 ** 30 if use_cuda else 0.0
            host_gb = psutil.Process().memory_info().rss / 2 ** 30
            eta_min = (total_epochs - epoch) * dt / 60.0
            print(f"[epoch {epoch}] PeakMemory(GPU) {gpu_gb:.2f} GiB | HostMemory {host_gb:.2f} GiB | "
                  f"estimated time left {eta_min:.1f} min", flush=True)
            tracker.log({**{f"loss/{k}": v for k, v in msg.items()}, "learning_rate": opt.param_groups[0]["lr"],
                         "throughput/images_per_sec": seen / dt, "memory/gpu_peak_gib": gpu_gb,
                         "memory/host_rss_gib": host_gb}, step)
        if epoch % eval_period == 0:
            model.eval()
            if val_set is not None:      # each rank predicts its shard of the validation split
                from b200flow.data.coco import collate_padded

                mine = list(range(rank, min(len(val_set), 8 * world), world)) or [0]
                vimgs, vt = collate_padded([val_set[j] for j in mine])
                val = [{"image": vimgs[j], **vt[j]} for j in range(len(mine))]
            else:
                val = list(coco_like(4, H, W, seed=77 + rank, device="cpu"))
            scores = evaluate_coco(model, val, dev, use_cuda, mcfg.mode_mask)
            if rank == 0:
                print(f"[epoch {epoch}] " + " | ".join(f"{k}: {v:.4f}" for k, v in scores.items()), flush=True)
                tracker.log(scores, step)
        if epoch % ckpt_period == 0 and rank == 0:
            name = f"model-{step}"
            torch.save({"model": model.state_dict(), "optimizer": opt.state_dict(), "global_step": step,
                        "config": cfgd, "model_config": mcfg.__dict__}, logdir / (name + ".pt"))
            (logdir / (name + ".index")).write_text(name + ".pt\n")
            ckfile.write_text(json.dumps({"model_checkpoint_path": name}))
        if world > 1:
            dist.barrier()
    if tracker:
        tracker.finish()
    if ddp is not None and ddp.timeline is not None:
        ddp.timeline.flush()
    if tuner is not None and tuner.done and rank == 0:
        print(f"[b200flow] HOROVOD_AUTOTUNE: bucket cap {tuner.best_mb} MB", flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
