"""BASELINE.json config 3: Mask R-CNN ResNet-50-FPN on COCO-shaped synthetic images, DDP — images/s.

Same trainer pieces as `workloads/maskrcnn/train.py` (bf16 channels_last backbone, K12-K15
detection kernels, flat momentum SGD, K1 bucketed all-reduce over peer memory for N > 1), images of
800x1344 with synthetic boxes / labels / masks, `TRAIN.BATCH_SIZE_PER_GPU` images per GPU (default 4,
the aws-samples values file: examples/maskrcnn/train-maskrcnn-aws.yaml).
"""
from __future__ import annotations

import json

import torch
import torch.distributed as dist


def run_maskrcnn_bench(args, rank, world, dev, ClockSampler, emit=None):
    from b200flow.data.synthetic import coco_like
    from b200flow.models.maskrcnn import MaskRCNN, MaskRCNNConfig
    from b200flow.ops import _lib
    from b200flow.optim import FlatMomentumSGD
    from b200flow.parallel.ddp import DistributedDataParallel
    from b200flow.utils.flat import FlatParams

    batch = args.micro_batch if args.micro_batch > 1 else 4
    H, W = 800, 1344
    torch.manual_seed(1234)
    model = MaskRCNN(MaskRCNNConfig()).to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    ddp = DistributedDataParallel(model, fused=not args.no_fused) if world > 1 else None
    flat = ddp.flat if ddp is not None else FlatParams(model)
    peak_lr = 0.01 * world * batch / 8.0
    opt = FlatMomentumSGD(flat, lr=peak_lr, momentum=0.9, weight_decay=1e-4)
    model.train()
    data = coco_like(10 ** 9, H, W, seed=1000 + rank, device="cpu")

    def make():
        samples = [next(data) for _ in range(batch)]
        return (torch.stack([s["image"] for s in samples]).to(torch.bfloat16).pin_memory(),
                [{k: s[k].pin_memory() for k in ("boxes", "labels", "masks")} for s in samples])

    host = [make() for _ in range(3)]
    import os

    graphed = False
    if model.static_shapes and os.environ.get("B2F_MASKRCNN_GRAPH", "1") == "1":
        graphed = model.enable_trunk_graph(host[0][0].to(dev))
    devb = [(im.to(dev), [{k: v.to(dev) for k, v in t.items()} for t in tg]) for im, tg in host]
    state = {"i": 0}

    def step(src, read_loss):
        im, tg = src[state["i"] % len(src)]
        state["i"] += 1
        if src is host:
            im = im.to(dev, non_blocking=True)
            tg = [{k: v.to(dev, non_blocking=True) for k, v in t.items()} for t in tg]
        # linear-scaling rule with a linear warm-up (tensorpack warms up as well): at 8 GPUs the full rate of 0.04
        # from a cold start diverges on a randomly initialised backbone
        opt.param_groups[0]["lr"] = peak_lr * min(1.0, (state["i"] + 1) / 100.0)
        opt.zero_grad()
        losses = model(im, tg)
        losses["total"].backward()
        flat.collect_autograd_grads()
        if ddp is not None:
            ddp.finish_reduction()
        opt.step()
        return float(losses["total"].detach()) if read_loss else losses["total"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    first = after_warmup = None
    for _ in range(args.warmup):
        after_warmup = step(devb, True)
        first = first if first is not None else after_warmup
    _lib.reset_launch_count()
    with ClockSampler(torch.cuda.current_device()) as clocks:
        ms = timed(lambda: step(devb, False), args.steps)
    launches = _lib.launch_count()
    last = step(host, True)
    ms_e2e = timed(lambda: step(host, True), args.steps)
    if rank == 0:
        gb = batch * world
        h2d = host[0][0].numel() * 2 + sum(v.numel() * v.element_size() for t in host[0][1] for v in t.values())
        (emit or (lambda o: print(json.dumps(o))))({
            "metric": "Mask R-CNN R50-FPN training images/s", "value": gb * args.steps / (ms / 1e3),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "b200flow",
            "config": {"model": "maskrcnn-r50-fpn", "global_batch": gb, "image": [H, W],
                       "parallelism": f"dp{world}", "l2_policy": "activations per step >> 126 MB L2",
                       "fused_collectives": not args.no_fused, "trunk_cuda_graph": graphed,
                       "lr": f"{peak_lr:g} (0.01 x global batch / 8), linear warm-up over 100 steps"},
            "clocks": clocks.summary(),
            "e2e": {"value": gb * args.steps / (ms_e2e / 1e3), "unit": "images/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
            "gpu_launches": launches, "loss_first_last": [first, last], "loss_after_warmup": after_warmup,
            **({} if (last == last and abs(last) < 1e4) else {"invalid": "training diverged: the loss is not finite"}),
            "native_libs": _lib.loaded_libraries()})
    if world > 1:
        dist.destroy_process_group()
    return 0
