"""BERT pre-training (masked LM + next-sentence prediction) through the Accelerate-shaped API.

BASELINE.json config 2 — "BERT-large pretraining bf16 Accelerate DDP" — as a program a user runs (bench.py
times the same step on synthetic batches):

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m b200flow.workloads.pretrain_bert \
        --model bert-large --data-prefix /fsx/data/bert/wiki_text_sentence --vocab-file vocab.txt \
        --max-seq-length 512 --per-device-batch 64 --max-steps 100000 --lr 1e-4 --warmup-steps 10000 \
        --mixed_precision bf16 --output_dir /fsx/home/bert/checkpoints --checkpointing_steps 5000 --with_tracking

`--data-prefix` names a sentence-level mmap corpus (`tools/preprocess_data.py --split-sentences`); samples come
from `data/bert_dataset.py` (C++ sentence-pair map, masking in the loader).  Without it the run uses random
tokens of the same shape.  Gradients are reduced by the bucketed K1 reducer during backward, parameters are
updated by the fused multi-tensor AdamW, checkpoints are Accelerate `save_state` directories (`step_<n>`) and
`--resume_from_checkpoint` continues from one, skipping the batches already seen.
"""
from __future__ import annotations

import argparse
import os
import time

import torch
from torch.utils.data import DataLoader

from b200flow.engines.accelerate_api import Accelerator
from b200flow.models.bert import BertForPreTraining, bert_config
from b200flow.obs.tracking import maybe_inject_fault
from b200flow.optim import AdamW, get_linear_schedule_with_warmup


class SyntheticPretraining(torch.utils.data.Dataset):
    """Random-token samples with the masking statistics of the real loader (no corpus mounted)."""

    def __init__(self, vocab_size: int, seq: int, n: int, seed: int = 0):
        self.vocab, self.seq, self.n, self.seed = vocab_size, seq, n, seed

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 7919 + i)
        ids = torch.randint(5, self.vocab, (self.seq,), generator=g)
        labels = torch.full((self.seq,), -100, dtype=torch.long)
        m = torch.rand(self.seq, generator=g) < 0.15
        m[0] = False
        labels[m] = ids[m]
        ids = ids.clone()
        ids[m] = 4
        half = self.seq // 2
        return {"input_ids": ids, "token_type_ids": (torch.arange(self.seq) >= half).long(),
                "attention_mask": torch.ones(self.seq, dtype=torch.long), "labels": labels,
                "next_sentence_label": torch.randint(0, 2, (), generator=g)}


def collate(rows):
    out = {k: torch.stack([r[k] for r in rows]) for k in rows[0]}
    # the collator hands the model the flat positions of the predicted tokens (no device-side search)
    out["masked_lm_positions"] = (out["labels"].reshape(-1) != -100).nonzero().squeeze(-1)
    return out


def parse_args(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--model", default="bert-large")
    p.add_argument("--data-prefix", default=None)
    p.add_argument("--vocab-file", default=None)
    p.add_argument("--max-seq-length", type=int, default=512)
    p.add_argument("--per-device-batch", type=int, default=64)
    p.add_argument("--gradient_accumulation_steps", type=int, default=1)
    p.add_argument("--max-steps", type=int, default=1000)
    p.add_argument("--lr", type=float, default=1e-4)
    p.add_argument("--weight-decay", type=float, default=0.01)
    p.add_argument("--warmup-steps", type=int, default=100)
    p.add_argument("--clip-grad", type=float, default=1.0)
    p.add_argument("--masked-lm-prob", type=float, default=0.15)
    p.add_argument("--short-seq-prob", type=float, default=0.1)
    p.add_argument("--dropout", type=float, default=0.1)
    p.add_argument("--seed", type=int, default=1234)
    p.add_argument("--mixed_precision", default="bf16", choices=["no", "fp16", "bf16"])
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--output_dir", default=None)
    p.add_argument("--checkpointing_steps", type=int, default=0)
    p.add_argument("--resume_from_checkpoint", default=None)
    p.add_argument("--with_tracking", action="store_true")
    p.add_argument("--project_dir", default="logs")
    p.add_argument("--log-interval", type=int, default=10)
    return p.parse_args(argv)


def main(argv=None) -> int:
    args = parse_args(argv)
    acc = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision,
                      gradient_accumulation_steps=args.gradient_accumulation_steps,
                      log_with="all" if args.with_tracking else None, project_dir=args.project_dir)
    over = {}
    special = dict(cls_id=2, sep_id=3, mask_id=4, pad_id=0)
    if args.vocab_file:
        from b200flow.data.tokenizer import BertWordPieceTokenizer

        tok = BertWordPieceTokenizer(args.vocab_file)
        over["vocab_size"] = (tok.vocab_size + 127) // 128 * 128
        special = dict(cls_id=tok.cls, sep_id=tok.sep, mask_id=tok.mask, pad_id=tok.pad)
    cfg = bert_config(args.model, max_pos=max(args.max_seq_length, 128), dropout=args.dropout, **over)
    torch.manual_seed(args.seed)
    model = BertForPreTraining(cfg, dtype=acc.param_dtype, device=acc.device)
    total = args.max_steps * args.per_device_batch * args.gradient_accumulation_steps * acc.num_processes
    if args.data_prefix:
        from b200flow.data.bert_dataset import BertPretrainingDataset

        with acc.main_process_first():
            data = BertPretrainingDataset(args.data_prefix, vocab_size=cfg.vocab_size, max_seq_length=args.max_seq_length,
                                          num_samples=total, masked_lm_prob=args.masked_lm_prob,
                                          short_seq_prob=args.short_seq_prob, seed=args.seed, **special)
    else:
        data = SyntheticPretraining(cfg.vocab_size, args.max_seq_length, total, args.seed)
    loader = DataLoader(data, batch_size=args.per_device_batch, shuffle=False, collate_fn=collate, drop_last=True)
    opt = AdamW(model.parameters(), lr=args.lr, betas=(0.9, 0.999), weight_decay=args.weight_decay)
    sched = get_linear_schedule_with_warmup(opt, args.warmup_steps, args.max_steps)
    model, opt, loader, sched = acc.prepare(model, opt, loader, sched)
    if args.with_tracking:
        acc.init_trackers("pretrain_bert", vars(args))
    step = 0
    if args.resume_from_checkpoint:
        acc.load_state(args.resume_from_checkpoint)
        step = int(os.path.basename(os.path.normpath(args.resume_from_checkpoint)).split("_")[-1])
        loader = acc.skip_first_batches(loader, step * args.gradient_accumulation_steps)
        acc.print(f"resumed from {args.resume_from_checkpoint} at step {step}")
    model.train()
    t0, seen, running = time.time(), 0, 0.0
    for batch in loader:
        with acc.accumulate(model):
            out = model(**batch)
            acc.backward(out["loss"])
            if acc.sync_gradients and args.clip_grad > 0:
                acc.clip_grad_norm_(model.parameters(), args.clip_grad)
            opt.step()
            sched.step()
            opt.zero_grad()
        running += float(out["loss"].detach())
        seen += 1
        if not acc.sync_gradients:
            continue
        step += 1
        maybe_inject_fault(step)
        if step % args.log_interval == 0 or step == args.max_steps:
            dt = time.time() - t0
            sps = seen * args.per_device_batch * acc.num_processes / dt
            acc.print(f"step {step:7d}/{args.max_steps} | loss {running / seen:.4f} | lr {sched.get_last_lr()[0]:.3e} | "
                      f"{sps:.1f} samples/s")
            if args.with_tracking:
                acc.log({"train_loss": running / seen, "lr": sched.get_last_lr()[0], "samples_per_second": sps}, step=step)
            t0, seen, running = time.time(), 0, 0.0
        if args.output_dir and args.checkpointing_steps and step % args.checkpointing_steps == 0:
            acc.save_state(os.path.join(args.output_dir, f"step_{step}"))
        if step >= args.max_steps:
            break
    if args.output_dir:
        acc.save_state(os.path.join(args.output_dir, f"step_{step}"))
    if args.with_tracking:
        acc.end_training()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
