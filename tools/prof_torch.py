"""Kernel time shares of one training step, from torch.profiler (CUPTI): which kernels a step spends its time in.
Shares only - a duration taken under a profiler is never reported as a benchmark value.
usage: python tools/prof_torch.py maskrcnn|maskrcnn_dyn|moe_layer [rows]"""
import os
import sys

import torch

sys.path.insert(0, "/root/repo")
what = sys.argv[1]
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 28
dev = torch.device("cuda", 0)

if what.startswith("maskrcnn"):
    os.environ["B2F_MASKRCNN_STATIC"] = "0" if what.endswith("dyn") else "1"
    from b200flow.data.synthetic import coco_like
    from b200flow.models.maskrcnn import MaskRCNN, MaskRCNNConfig
    from b200flow.optim import FlatMomentumSGD
    from b200flow.utils.flat import FlatParams

    torch.manual_seed(1234)
    model = MaskRCNN(MaskRCNNConfig()).to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    flat = FlatParams(model)
    opt = FlatMomentumSGD(flat, lr=0.005, momentum=0.9, weight_decay=1e-4)
    model.train()
    data = coco_like(10 ** 9, 800, 1344, seed=1000, device="cpu")
    s = [next(data) for _ in range(4)]
    im = torch.stack([x["image"] for x in s]).to(torch.bfloat16).to(dev)
    tg = [{k: x[k].to(dev) for k in ("boxes", "labels", "masks")} for x in s]

    def step():
        opt.zero_grad()
        losses = model(im, tg)
        losses["total"].backward()
        flat.collect_autograd_grads()
        opt.step()
        return losses["total"]
else:
    from b200flow.parallel.moe import MoELayer

    layer = MoELayer(2048, 8192, 8, dtype=torch.bfloat16, device=dev, fp8=True)
    x = (torch.randn(16384, 2048, device=dev) * 0.5).to(torch.bfloat16).requires_grad_()

    def step():
        y = layer(x)
        y.float().sum().backward()
        return y

for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(4):
    out = step()
e1.record()
torch.cuda.synchronize()
print(f"# {what}: {e0.elapsed_time(e1) / 4:.2f} ms/step without the profiler; last value {float(out.float().sum().detach()):.4f}")
from torch.profiler import ProfilerActivity, profile

with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_type == torch.autograd.DeviceType.CUDA]
tot = sum(e.self_device_time_total for e in ev)
print(f"# device time {tot / 2e3:.2f} ms/step in {sum(e.count for e in ev) // 2} kernels/step")
for e in sorted(ev, key=lambda e: -e.self_device_time_total)[:rows]:
    print(f"{100 * e.self_device_time_total / tot:5.1f}%  {e.self_device_time_total / 2e3:8.3f} ms  n={e.count // 2:4d}  {e.key[:110]}")
