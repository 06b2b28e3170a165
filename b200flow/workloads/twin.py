"""Stock-library twin of the headline workloads (`bench.py --impl twin`): the on-box comparison arm.

The reference itself has nothing to install, so "what a user of the reference's stack
gets on this box" is approximated with the libraries that are in the image and nothing of ours:
plain `torch.nn` modules, cuBLAS through `F.linear`, `F.scaled_dot_product_attention` (cuDNN / flash),
`F.rms_norm` / `F.layer_norm`, `torch.optim.AdamW(fused=True)`, torch `DistributedDataParallel` over
NCCL for N > 1, and Hugging Face `BertForPreTraining` (sdpa attention) for BERT.  Same shapes, same
synthetic data, same timing protocol and JSON line as the b200flow arm; `"impl": "twin"`.

Differences that cannot be avoided are stated in `config.notes`: the Llama twin scales out with
data parallelism only (torch has no TP/PP engine to call), so its N > 1 points are DDP with a
replicated 7B model, not TP2·PP2·DP2.
"""
from __future__ import annotations

import json

import torch
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
        return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), labels.view(-1))


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


def run_twin(args, rank, world, dev, ClockSampler):
    torch.manual_seed(1234)
    is_bert = args.workload.startswith("bert")
    notes = []
    if is_bert:
        from transformers import BertConfig, BertForPreTraining

        seq = min(args.seq, 512)
        batch = args.micro_batch if args.micro_batch > 1 else 64
        cfg = BertConfig(vocab_size=30592, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
                         intermediate_size=4096, max_position_embeddings=512, attn_implementation="sdpa")
        model = BertForPreTraining(cfg).to(dev, torch.bfloat16)
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
            return mod(**b).loss
        metric, unit, per_step = "BERT-large pretraining samples/s", "samples/s", batch
        name = "bert-large"
        notes.append("HF BertForPreTraining(sdpa): decoder over all positions (HF has no masked-position gather)")
    else:
        seq, batch, gas = args.seq, args.micro_batch, args.gas
        model = TwinLlama(32000, 4096, 32, 32, 11008, seq, torch.bfloat16, dev)
        gen = torch.Generator().manual_seed(77 + rank)

        def make():
            t = torch.randint(0, 32000, (batch, seq + 1), generator=gen)
            return {"tokens": t[:, :-1].contiguous(), "labels": t[:, 1:].contiguous()}

        def loss_fn(mod, b):
            return mod(b["tokens"], b["labels"])
        metric, unit, per_step = "Llama-2 7B Megatron-DS tokens/s", "tokens/s", batch * seq * gas
        name = "llama2-7b"
        if world > 1:
            notes.append("torch DDP with a replicated model (stock torch has no TP/PP engine); not TP2.PP2.DP2")
    model.train()
    ddp = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index]) if world > 1 else model
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

    for _ in range(args.warmup):
        step(devb, True)
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
            "gpu_launches": 0}))
    if world > 1:
        dist.destroy_process_group()
    return 0
