"""Stock-library twin of the headline workloads (`bench.py --impl twin`): the on-box comparison arm.

The reference itself has nothing to install, so "what a user of the reference's stack
gets on this box" is approximated with the libraries that are in the image and nothing of ours:
plain `torch.nn` modules, cuBLAS through `F.linear`, `F.scaled_dot_product_attention` (cuDNN / flash),
`F.rms_norm` / `F.layer_norm`, `torch.optim.AdamW(fused=True)`, torch `DistributedDataParallel` over
NCCL for N > 1, Hugging Face `BertForPreTraining` (sdpa attention) for BERT, torchvision's
`maskrcnn_resnet50_fpn` for Mask R-CNN, and a DeepSpeed-MoE-recipe GPT (`all_to_all_single` + `F.linear`
experts) for GPT-MoE.  Same shapes, same synthetic data, same timing protocol and JSON line as the
b200flow arm; `"impl": "twin"`.

Differences that cannot be avoided are stated in `config.notes`: the Llama twin scales out with
data parallelism only (torch has no TP/PP engine to call), so its N > 1 points are DDP with a
replicated 7B model, not TP2·PP2·DP2.
"""
from __future__ import annotations

import json

import torch

from ..data.synthetic import zipf_tokens
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn


class _TwinLlamaLayer(nn.Module):
    def __init__(self, hidden, heads, ffn, dtype, device):
        super().__init__()
        kw = dict(bias=False, dtype=dtype, device=device)
        self.heads, self.hd = heads, hidden // heads
        self.qkv = nn.Linear(hidden, 3 * hidden, **kw)
        self.o = nn.Linear(hidden, hidden, **kw)
        self.gate_up = nn.Linear(hidden, 2 * ffn, **kw)
        self.down = nn.Linear(ffn, hidden, **kw)
        self.n1 = nn.Parameter(torch.ones(hidden, dtype=dtype, device=device))
        self.n2 = nn.Parameter(torch.ones(hidden, dtype=dtype, device=device))

    def forward(self, x, cos, sin):
        b, s, H = x.shape
        h = F.rms_norm(x, (H,), self.n1, 1e-5)
        q, k, v = self.qkv(h).view(b, s, 3, self.heads, self.hd).unbind(2)
        q, k = _rope(q, cos, sin), _rope(k, cos, sin)
        a = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
        x = x + self.o(a.transpose(1, 2).reshape(b, s, H))
        h = F.rms_norm(x, (H,), self.n2, 1e-5)
        g, u = self.gate_up(h).chunk(2, -1)
        return x + self.down(F.silu(g) * u)


def _rope(x, cos, sin):
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], -1)


class TwinLlama(nn.Module):
    def __init__(self, vocab, hidden, layers, heads, ffn, seq, dtype, device):
        super().__init__()
        self.emb = nn.Embedding(vocab, hidden, dtype=dtype, device=device)
        self.layers = nn.ModuleList([_TwinLlamaLayer(hidden, heads, ffn, dtype, device) for _ in range(layers)])
        self.norm = nn.Parameter(torch.ones(hidden, dtype=dtype, device=device))
        self.head = nn.Linear(hidden, vocab, bias=False, dtype=dtype, device=device)
        hd = hidden // heads
        inv = 1.0 / (10000 ** (torch.arange(0, hd, 2, device=device).float() / hd))
        ang = torch.outer(torch.arange(seq, device=device).float(), inv)
        self.register_buffer("cos", ang.cos().to(dtype)[None, :, None, :], persistent=False)
        self.register_buffer("sin", ang.sin().to(dtype)[None, :, None, :], persistent=False)
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.normal_(p, std=0.02)

    def forward(self, tokens, labels):
        x = self.emb(tokens)
        for layer in self.layers:
            x = layer(x, self.cos, self.sin)
        logits = self.head(F.rms_norm(x, (x.shape[-1],), self.norm, 1e-5))
        return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), labels.reshape(-1))


def _timed(fn, k, world, dev):
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
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


class _TwinMoE(nn.Module):
    """DeepSpeed-MoE's recipe with stock ops: fp32 gate -> softmax -> top-1 -> capacity cut by position in the
    expert's queue -> [E, C, H] dispatch buffer -> `all_to_all_single` -> per-expert `F.linear` FFN -> all-to-all
    back -> gate-weighted gather.  Index-based dispatch instead of DeepSpeed's dense [T, E, C] einsum masks, which
    only makes this arm faster."""

    def __init__(self, hidden, ffn, experts, world, group, dtype, device):
        super().__init__()
        self.E, self.El, self.world, self.group = experts, experts // world, world, group
        self.gate = nn.Linear(hidden, experts, bias=False, dtype=torch.float32, device=device)
        self.w1 = nn.Parameter(torch.empty(self.El, ffn, hidden, dtype=dtype, device=device).normal_(0, 0.02))
        self.b1 = nn.Parameter(torch.zeros(self.El, ffn, dtype=dtype, device=device))
        self.w2 = nn.Parameter(torch.empty(self.El, hidden, ffn, dtype=dtype, device=device).normal_(0, 0.02))
        self.b2 = nn.Parameter(torch.zeros(self.El, hidden, dtype=dtype, device=device))
        self.aux = None

    def forward(self, x):
        T, H = x.shape
        C = (max(4, -(-T // self.E)) + 7) // 8 * 8
        probs = self.gate(x.float()).softmax(-1)
        gate, best = probs.max(-1)
        onehot = F.one_hot(best, self.E)
        slot = (onehot.cumsum(0) - 1).gather(1, best[:, None]).squeeze(1)
        keep = slot < C
        self.aux = (probs.mean(0) * onehot.float().mean(0)).sum() * self.E
        rows = (best * C + slot).clamp(max=self.E * C - 1)
        buf = x.new_zeros(self.E * C, H)
        buf[rows[keep]] = x[keep]
        if self.world > 1:
            recv = torch.empty_like(buf)
            dist.all_to_all_single(recv, buf, group=self.group)
            buf = recv                                                    # [world(src), El, C, H]
        xin = buf.view(self.world, self.El, C, H).transpose(0, 1).reshape(self.El, self.world * C, H)
        out = torch.stack([F.linear(F.gelu(F.linear(xin[e], self.w1[e], self.b1[e]), approximate="tanh"),
                                    self.w2[e], self.b2[e]) for e in range(self.El)])
        back = out.view(self.El, self.world, C, H).transpose(0, 1).reshape(self.E * C, H)
        if self.world > 1:
            recv = torch.empty_like(back)
            dist.all_to_all_single(recv, back.contiguous(), group=self.group)
            back = recv
        y = back[rows] * (gate * keep)[:, None].to(x.dtype)
        return y


class _TwinGPTLayer(nn.Module):
    def __init__(self, hidden, heads, ffn, dtype, device, moe=None):
        super().__init__()
        kw = dict(dtype=dtype, device=device)
        self.heads, self.hd = heads, hidden // heads
        self.ln1, self.ln2 = nn.LayerNorm(hidden, **kw), nn.LayerNorm(hidden, **kw)
        self.qkv, self.o = nn.Linear(hidden, 3 * hidden, **kw), nn.Linear(hidden, hidden, **kw)
        self.moe = moe
        if moe is None:
            self.fc1, self.fc2 = nn.Linear(hidden, ffn, **kw), nn.Linear(ffn, hidden, **kw)

    def forward(self, x):
        b, s, H = x.shape
        q, k, v = self.qkv(self.ln1(x)).view(b, s, 3, self.heads, self.hd).unbind(2)
        a = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
        x = x + self.o(a.transpose(1, 2).reshape(b, s, H))
        h = self.ln2(x)
        if self.moe is not None:
            return x + self.moe(h.reshape(b * s, H)).view(b, s, H)
        return x + self.fc2(F.gelu(self.fc1(h), approximate="tanh"))


class TwinGPTMoE(nn.Module):
    """GPT with an MoE FFN in every `moe_every`-th layer (BASELINE.json config 5), stock torch modules only."""

    def __init__(self, vocab, hidden, layers, heads, ffn, seq, experts, moe_every, world, group, dtype, device,
                 aux_coeff: float = 0.01):
        super().__init__()
        self.emb = nn.Embedding(vocab, hidden, dtype=dtype, device=device)
        self.pos = nn.Embedding(seq, hidden, dtype=dtype, device=device)
        self.layers = nn.ModuleList([
            _TwinGPTLayer(hidden, heads, ffn, dtype, device,
                          _TwinMoE(hidden, ffn, experts, world, group, dtype, device) if (i + 1) % moe_every == 0 else None)
            for i in range(layers)])
        self.ln = nn.LayerNorm(hidden, dtype=dtype, device=device)
        self.aux_coeff = aux_coeff
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.normal_(p, std=0.02)

    def expert_parameters(self):
        return [p for l in self.layers if l.moe is not None for p in (l.moe.w1, l.moe.b1, l.moe.w2, l.moe.b2)]

    def forward(self, tokens, labels):
        x = self.emb(tokens) + self.pos(torch.arange(tokens.shape[1], device=tokens.device))[None]
        for layer in self.layers:
            x = layer(x)
        logits = F.linear(self.ln(x), self.emb.weight)
        loss = F.cross_entropy(logits.float().view(-1, logits.shape[-1]), labels.reshape(-1))
        aux = sum(l.moe.aux for l in self.layers if l.moe is not None)
        return loss + self.aux_coeff * aux


def maskrcnn_twin(dev, dtype, batch: int, H: int, W: int, seed: int):
    """torchvision's Mask R-CNN R50-FPN (its own RoIAlign / NMS / box-coder CUDA ops, cuDNN convolutions) on the
    same COCO-shaped synthetic stream as `bench.py --workload maskrcnn`.  Returns (model, make_batch, loss_fn)."""
    import torchvision

    from ..data.synthetic import coco_like

    model = torchvision.models.detection.maskrcnn_resnet50_fpn(
        weights=None, weights_backbone=None, num_classes=81, min_size=min(H, W), max_size=max(H, W),
        image_mean=[0.0, 0.0, 0.0], image_std=[1.0, 1.0, 1.0]).to(dev)
    stream = coco_like(10 ** 9, H, W, seed=seed, device="cpu")

    def make():
        samples = [next(stream) for _ in range(batch)]
        out = {"images": torch.stack([s["image"] for s in samples])}
        for i, s in enumerate(samples):          # flat dict of tensors so that pin / copy loops stay generic
            out[f"boxes{i}"], out[f"labels{i}"], out[f"masks{i}"] = s["boxes"], s["labels"], s["masks"]
        return out

    def loss_fn(mod, b):
        n = b["images"].shape[0]
        targets = [{"boxes": b[f"boxes{i}"], "labels": b[f"labels{i}"], "masks": b[f"masks{i}"]} for i in range(n)]
        with torch.autocast(device_type=b["images"].device.type, dtype=dtype, enabled=dtype != torch.float32):
            losses = mod(list(b["images"]), targets)
        return sum(losses.values())

    return model, make, loss_fn


def run_twin(args, rank, world, dev, ClockSampler):
    torch.manual_seed(1234)
    is_bert = args.workload.startswith("bert")
    is_mrcnn = args.workload.startswith("maskrcnn")
    notes = []
    if is_mrcnn:
        batch = args.micro_batch if args.micro_batch > 1 else 4
        seq, gas = 0, 1
        model, make, loss_fn = maskrcnn_twin(dev, torch.bfloat16, batch, 800, 1344, 1000 + rank)
        metric, unit, per_step, name = "Mask R-CNN R50-FPN images/s", "images/s", batch, "maskrcnn-r50-fpn"
        notes.append("torchvision maskrcnn_resnet50_fpn under bf16 autocast, SGD momentum (torch.optim.SGD)")
    elif "moe" in args.workload:
        seq = 2048 if args.seq == 4096 else args.seq
        batch, gas = (8, 4) if (args.micro_batch == 1 and args.gas == 16) else (args.micro_batch, args.gas)
        ep = min(world, 8)
        model = TwinGPTMoE(50304, 2048, 24, 16, 8192, seq, 8, 2, ep, dist.group.WORLD if world > 1 else None,
                           torch.bfloat16, dev)
        gen = torch.Generator().manual_seed(77 + rank)

        def make():
            t = zipf_tokens(50304, (batch, seq + 1), generator=gen)
            return {"tokens": t[:, :-1].contiguous(), "labels": t[:, 1:].contiguous()}

        def loss_fn(mod, b):
            return mod(b["tokens"], b["labels"])
        metric, unit, per_step, name = "GPT-MoE 1.3B (8 experts) tokens/s", "tokens/s", batch * seq * gas, "gpt-moe-1.3b"
        notes.append("bf16 experts through F.linear (stock torch has no fp8 training GEMM path with scaling here); "
                     "expert gradients are excluded from the DDP all-reduce (each rank owns its experts)")
    elif is_bert:
        from transformers import BertConfig, BertForPreTraining

        seq = min(args.seq, 512)
        batch = args.micro_batch if args.micro_batch > 1 else 64
        cfg = BertConfig(vocab_size=30592, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
                         intermediate_size=4096, max_position_embeddings=512, attn_implementation="sdpa")
        hf = BertForPreTraining(cfg).to(dev, torch.bfloat16)

        class _GatheredHeads(nn.Module):
            """HF encoder + heads with the MLM decoder applied to the masked positions only - the same work the
            b200flow model does (NVIDIA-style `masked_lm_positions`), so the two arms are like for like."""

            def __init__(self, m):
                super().__init__()
                self.m = m

            def forward(self, input_ids, token_type_ids, attention_mask, labels, next_sentence_label):
                o = self.m.bert(input_ids=input_ids, token_type_ids=token_type_ids, attention_mask=attention_mask)
                seq_out, pooled = o[0], o[1]
                flat = labels.reshape(-1)
                pos = (flat != -100).nonzero().squeeze(-1)
                logits = self.m.cls.predictions(seq_out.reshape(-1, seq_out.shape[-1])[pos])
                mlm = F.cross_entropy(logits.float(), flat[pos])
                nsp = F.cross_entropy(self.m.cls.seq_relationship(pooled).float(), next_sentence_label)
                return mlm + nsp

        model = _GatheredHeads(hf)
        gas = 1
        gen = torch.Generator().manual_seed(77 + rank)

        def make():
            ids = torch.randint(1000, cfg.vocab_size - 200, (batch, seq), generator=gen)
            labels = torch.full((batch, seq), -100, dtype=torch.long)
            m = torch.rand(batch, seq, generator=gen) < 0.15
            labels[m] = ids[m]
            ids = ids.clone()
            ids[m] = 103
            return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": torch.ones_like(ids),
                    "labels": labels, "next_sentence_label": torch.randint(0, 2, (batch,), generator=gen)}

        def loss_fn(mod, b):
            return mod(**b)
        metric, unit, per_step = "BERT-large pretraining samples/s", "samples/s", batch
        name = "bert-large"
        notes.append("HF BertForPreTraining(sdpa) encoder and heads; MLM decoder on the masked positions only, as in the b200flow arm")
    else:
        seq, batch, gas = args.seq, args.micro_batch, args.gas
        model = TwinLlama(32000, 4096, 32, 32, 11008, seq, torch.bfloat16, dev)
        gen = torch.Generator().manual_seed(77 + rank)

        def make():
            t = torch.randint(0, 32000, (batch, seq + 1), generator=gen)       # as bench.py's Llama arm
            return {"tokens": t[:, :-1].contiguous(), "labels": t[:, 1:].contiguous()}

        def loss_fn(mod, b):
            return mod(b["tokens"], b["labels"])
        metric, unit, per_step = "Llama-2 7B Megatron-DS tokens/s", "tokens/s", batch * seq * gas
        name = "llama2-7b"
        if world > 1:
            notes.append("torch DDP with a replicated model (stock torch has no TP/PP engine); not TP2.PP2.DP2")
    model.train()
    if world > 1 and hasattr(model, "expert_parameters"):      # experts differ per rank: keep them out of DDP
        names = {id(p) for p in model.expert_parameters()}
        nn.parallel.DistributedDataParallel._set_params_and_buffers_to_ignore_for_model(
            model, [n for n, p in model.named_parameters() if id(p) in names])
    ddp = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index],
                                              find_unused_parameters=is_mrcnn) if world > 1 else model
    if is_mrcnn:
        opt = torch.optim.SGD([p for p in model.parameters() if p.requires_grad], lr=0.01, momentum=0.9, weight_decay=1e-4)
    else:
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4, betas=(0.9, 0.95), weight_decay=0.01, fused=True)
    host = [[{k: v.pin_memory() for k, v in make().items()} for _ in range(gas)] for _ in range(2)]
    devb = [[{k: v.to(dev) for k, v in mb.items()} for mb in step] for step in host]
    state = {"i": 0}

    def step(src, read_loss):
        mbs = src[state["i"] % len(src)]
        state["i"] += 1
        loss = None
        for j, mb in enumerate(mbs):
            if src is host:
                mb = {k: v.to(dev, non_blocking=True) for k, v in mb.items()}
            if world > 1 and j + 1 < len(mbs):
                with ddp.no_sync():
                    loss = loss_fn(ddp, mb) / len(mbs)
                    loss.backward()
            else:
                loss = loss_fn(ddp, mb) / len(mbs)
                loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        return float(loss) if read_loss else loss

    warm_loss = None
    for _ in range(args.warmup):
        warm_loss = step(devb, True)
    with ClockSampler(torch.cuda.current_device()) as clocks:
        ms = _timed(lambda: step(devb, False), args.steps, world, dev)
    ms_e2e = _timed(lambda: step(host, True), args.steps, world, dev)
    if rank == 0:
        total = per_step * world
        print(json.dumps({
            "metric": metric, "value": total * args.steps / (ms / 1e3), "unit": unit, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "impl": "twin",
            "config": {"model": name, "global_batch": batch * gas * world, "seq_len": seq,
                       "parallelism": f"dp{world}", "stack": "torch nn + cuBLAS + SDPA + fused AdamW + torch DDP/NCCL",
                       "notes": notes},
            "clocks": clocks.summary(),
            "e2e": {"value": total * args.steps / (ms_e2e / 1e3), "unit": unit,
                    "h2d_bytes_per_step": sum(v.numel() * v.element_size() for mb in host[0] for v in mb.values()),
                    "d2h_bytes_per_step": 4},
            "gpu_launches": 0, "loss_after_warmup": warm_loss}))
    if world > 1:
        dist.destroy_process_group()
    return 0
