"""Drop-in for Accelerate's `examples/complete_nlp_example.py` on the b200flow runtime.

The reference runs the upstream script unchanged
(examples/accelerate/bert-glue-mrpc/pretrain.yaml:41-51:
`torchrun $DISTRIBUTED_ARGS examples/complete_nlp_example.py --mixed_precision fp16
--checkpointing_steps epoch --with_tracking --output_dir $CKPT_DIR --project_dir $PROJECT_DIR`).
Same flags, same training recipe ([UP] constants in SURVEY.md §2D W1: bert-base-cased,
lr 2e-5, 3 epochs, seed 42, batch 16 / eval 32, AdamW, linear schedule with 100 warm-up steps,
accuracy + F1 per epoch, `epoch_<n>` / `step_<n>` checkpoints, resume).  Data is MRPC-shaped
synthetic (no network); weights are random-init.

Size knobs for smoke tests (environment): B2F_MODEL (bert-base|bert-tiny), B2F_EPOCHS,
B2F_TRAIN_SAMPLES, B2F_EVAL_SAMPLES.
"""
import argparse
import os

import torch
from torch.utils.data import DataLoader

from b200flow.data.synthetic import collate_pad, mrpc_like
from b200flow.engines.accelerate_api import Accelerator
from b200flow.models.bert import BertForSequenceClassification, bert_config
from b200flow.obs.tracking import maybe_inject_fault
from b200flow.optim import AdamW, get_linear_schedule_with_warmup

MAX_GPU_BATCH_SIZE = 16
EVAL_BATCH_SIZE = 32


def f1_and_accuracy(preds: torch.Tensor, refs: torch.Tensor) -> dict:
    preds, refs = preds.long().cpu(), refs.long().cpu()
    tp = int(((preds == 1) & (refs == 1)).sum())
    fp = int(((preds == 1) & (refs == 0)).sum())
    fn = int(((preds == 0) & (refs == 1)).sum())
    f1 = 2 * tp / max(1, 2 * tp + fp + fn)
    return {"accuracy": float((preds == refs).float().mean()), "f1": f1}


def get_dataloaders(accelerator: Accelerator, batch_size: int, vocab: int):
    with accelerator.main_process_first():
        train = mrpc_like("train", vocab=vocab)
        val = mrpc_like("validation", vocab=vocab)
    nt, nv = os.environ.get("B2F_TRAIN_SAMPLES"), os.environ.get("B2F_EVAL_SAMPLES")
    if nt:
        train = train[:int(nt)]
    if nv:
        val = val[:int(nv)]
    g = torch.Generator().manual_seed(42)
    train_dl = DataLoader(train, shuffle=True, collate_fn=collate_pad, batch_size=batch_size, generator=g)
    eval_dl = DataLoader(val, shuffle=False, collate_fn=collate_pad, batch_size=EVAL_BATCH_SIZE)
    return train_dl, eval_dl


def training_function(config, args):
    if args.with_tracking:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, log_with="all",
                                  project_dir=args.project_dir)
    else:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    checkpointing_steps = args.checkpointing_steps
    if checkpointing_steps is not None and checkpointing_steps != "epoch":
        checkpointing_steps = int(checkpointing_steps)
    lr, num_epochs, seed, batch_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"])
    if args.with_tracking:
        accelerator.init_trackers(os.path.split(__file__)[-1].split(".")[0], config)
    gradient_accumulation_steps = 1
    if batch_size > MAX_GPU_BATCH_SIZE:
        gradient_accumulation_steps = batch_size // MAX_GPU_BATCH_SIZE
        batch_size = MAX_GPU_BATCH_SIZE
    accelerator.gradient_accumulation_steps = gradient_accumulation_steps
    torch.manual_seed(seed)
    cfg = bert_config(os.environ.get("B2F_MODEL", "bert-base"), num_labels=2)
    train_dl, eval_dl = get_dataloaders(accelerator, batch_size, cfg.vocab_size)
    model = BertForSequenceClassification(cfg, dtype=torch.float32)
    optimizer = AdamW(params=model.parameters(), lr=lr)
    lr_scheduler = get_linear_schedule_with_warmup(
        optimizer, num_warmup_steps=100,
        num_training_steps=(len(train_dl) * num_epochs) // gradient_accumulation_steps)
    model, optimizer, train_dl, eval_dl, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dl, eval_dl, lr_scheduler)

    overall_step, starting_epoch, resume_step = 0, 0, None
    if not args.resume_from_checkpoint and os.environ.get("B2F_AUTO_RESUME") == "1" and args.output_dir and \
            os.path.isdir(args.output_dir) and any(d.is_dir() and d.name.startswith(("epoch_", "step_"))
                                                   for d in os.scandir(args.output_dir)):
        args.resume_from_checkpoint = "latest"     # launcher --auto-resume
    if args.resume_from_checkpoint:
        path = args.resume_from_checkpoint
        if path in ("", "latest", True):
            dirs = sorted((d for d in os.scandir(args.output_dir or os.getcwd()) if d.is_dir()),
                          key=lambda d: os.path.getctime(d.path))
            path = dirs[-1].path
        accelerator.print(f"Resumed from checkpoint: {path}")
        accelerator.load_state(path)
        tag = os.path.basename(os.path.normpath(path))
        if "epoch" in tag:
            starting_epoch = int(tag.replace("epoch_", "")) + 1
        else:
            resume_step = int(tag.replace("step_", ""))
            starting_epoch = resume_step // len(train_dl)
            resume_step -= starting_epoch * len(train_dl)

    for epoch in range(starting_epoch, num_epochs):
        model.train()
        total_loss = 0.0
        if args.resume_from_checkpoint and epoch == starting_epoch and resume_step is not None:
            active = accelerator.skip_first_batches(train_dl, resume_step)
            overall_step += resume_step
        else:
            active = train_dl
        for step, batch in enumerate(active):
            maybe_inject_fault(overall_step)
            with accelerator.accumulate(model):
                outputs = model(**batch)
                loss = outputs["loss"]
                total_loss += float(loss.detach())
                accelerator.backward(loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
            overall_step += 1
            if isinstance(checkpointing_steps, int) and overall_step % checkpointing_steps == 0:
                out = f"step_{overall_step}"
                accelerator.save_state(os.path.join(args.output_dir, out) if args.output_dir else out)
        model.eval()
        preds, refs = [], []
        for step, batch in enumerate(eval_dl):
            with torch.no_grad():
                outputs = model(**batch)
            p = outputs["logits"].argmax(dim=-1)
            p, r = accelerator.gather_for_metrics((p, batch["labels"]))
            preds.append(p)
            refs.append(r)
        metric = f1_and_accuracy(torch.cat(preds), torch.cat(refs))
        accelerator.print(f"epoch {epoch}:", metric)
        if args.with_tracking:
            accelerator.log({"accuracy": metric["accuracy"], "f1": metric["f1"],
                             "train_loss": total_loss / max(1, len(train_dl)), "epoch": epoch}, step=epoch)
        if checkpointing_steps == "epoch":
            out = f"epoch_{epoch}"
            accelerator.save_state(os.path.join(args.output_dir, out) if args.output_dir else out)
    accelerator.end_training()


def main():
    p = argparse.ArgumentParser(description="Simple example of training script.")
    p.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16", "fp8"])
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--checkpointing_steps", type=str, default=None)
    p.add_argument("--resume_from_checkpoint", type=str, default=None)
    p.add_argument("--with_tracking", action="store_true")
    p.add_argument("--output_dir", type=str, default=".")
    p.add_argument("--project_dir", type=str, default="logs")
    args, unknown = p.parse_known_args()
    if unknown:
        print(f"[b200flow] ignoring unknown flags: {unknown}")
    config = {"lr": 2e-5, "num_epochs": int(os.environ.get("B2F_EPOCHS", 3)), "seed": 42, "batch_size": 16}
    training_function(config, args)


if __name__ == "__main__":
    main()
