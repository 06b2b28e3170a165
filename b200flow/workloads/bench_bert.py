"""BASELINE.json config 2: BERT-large pre-training (MLM + NSP), bf16, DDP — samples/s.

Public-API path: `Accelerator.prepare` (bucketed K1 DDP over NVLink peer memory, fused AdamW),
synthetic token batches of the named shape (seq 512, 15 % masked), random-init weights.
"""
from __future__ import annotations

import json

import torch
import torch.distributed as dist


def _batches(cfg, n, batch, seq, gen):
    out = []
    for _ in range(n):
        ids = torch.randint(1000, cfg.vocab_size - 200, (batch, seq), generator=gen)
        labels = torch.full((batch, seq), -100, dtype=torch.long)
        m = torch.rand(batch, seq, generator=gen) < 0.15
        labels[m] = ids[m]
        ids = ids.clone()
        ids[m] = 103
        out.append({"input_ids": ids, "token_type_ids": torch.zeros_like(ids),
                    "attention_mask": torch.ones_like(ids), "labels": labels,
                    "next_sentence_label": torch.randint(0, 2, (batch,), generator=gen),
                    # collator-side list of the masked positions (NVIDIA-style masked_lm_positions)
                    "masked_lm_positions": (labels.reshape(-1) != -100).nonzero().squeeze(-1)})
    return out


def run_bert_bench(args, rank, world, dev, ClockSampler, emit=None):
    from b200flow.engines.accelerate_api import Accelerator
    from b200flow.models.bert import BertForPreTraining, bert_config
    from b200flow.ops import _lib
    from b200flow.optim import AdamW

    name = "bert-large" if args.workload in ("bert", "bert-large") else args.workload
    cfg = bert_config(name, dropout=0.1)
    seq = min(args.seq, cfg.max_pos)
    batch = args.micro_batch if args.micro_batch > 1 else 64
    acc = Accelerator(mixed_precision="bf16", fused_collectives=not args.no_fused)
    torch.manual_seed(1234)
    model = BertForPreTraining(cfg, dtype=torch.bfloat16, device=dev)
    opt = AdamW(model.parameters(), lr=1e-4, betas=(0.9, 0.999), weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    model.train()
    gen = torch.Generator().manual_seed(77 + rank)
    host = [{k: v.pin_memory() for k, v in b.items()} for b in _batches(cfg, 4, batch, seq, gen)]
    devb = [{k: v.to(dev) for k, v in b.items()} for b in host]
    state = {"i": 0}

    def step(src, read_loss):
        b = src[state["i"] % len(src)]
        state["i"] += 1
        if src is host:
            b = {k: v.to(dev, non_blocking=True) for k, v in b.items()}
        out = model(**b)
        acc.backward(out["loss"])
        opt.step()
        opt.zero_grad()
        return float(out["loss"]) if read_loss else out["loss"]

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

    first = None
    for _ in range(args.warmup):
        l = step(devb, True)
        first = first if first is not None else l
    _lib.reset_launch_count()
    with ClockSampler(torch.cuda.current_device()) as clocks:
        ms = timed(lambda: step(devb, False), args.steps)
    launches = _lib.launch_count()
    last = step(host, True)
    ms_e2e = timed(lambda: step(host, True), args.steps)
    if rank == 0:
        gb = batch * world
        (emit or (lambda o: print(json.dumps(o))))({
            "metric": "BERT-large pretraining samples/s", "value": gb * args.steps / (ms / 1e3),
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "b200flow",
            "config": {"model": name, "global_batch": gb, "seq_len": seq, "parallelism": f"dp{world}",
                       "l2_policy": "weights + activations per step >> 126 MB L2",
                       "fused_collectives": not args.no_fused, "params": cfg.param_count()},
            "clocks": clocks.summary(),
            "e2e": {"value": gb * args.steps / (ms_e2e / 1e3), "unit": "samples/s",
                    "h2d_bytes_per_step": sum(v.numel() * v.element_size() for v in host[0].values()),
                    "d2h_bytes_per_step": 4},
            "gpu_launches": launches, "loss_first_last": [first, last],
            "native_libs": _lib.loaded_libraries()})
    if world > 1:
        dist.destroy_process_group()
    return 0
