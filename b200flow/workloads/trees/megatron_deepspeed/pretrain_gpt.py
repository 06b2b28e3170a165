"""Drop-in for Megatron-DeepSpeed `pretrain_gpt.py` on the b200flow runtime.

Accepts the flags the reference passes (examples/megatron-deepspeed/gpt2_345m/
pretrain-ddp-zero1.yaml:39-55,76-84 and pretrain-ddp-tp-pp-zero1.yaml:39-57):
--num-layers --hidden-size --num-attention-heads --seq-length --max-position-embeddings
--micro-batch-size --global-batch-size --lr --train-iters --lr-decay-iters --lr-decay-style
--min-lr --weight-decay --lr-warmup-fraction --clip-grad --fp16/--bf16
--tensor-model-parallel-size --pipeline-model-parallel-size --sequence-parallel
--data-path --vocab-file --merge-file --data-impl --split --log-interval --save-interval
--eval-interval --eval-iters --distributed-backend --save --load --deepspeed
--deepspeed_config <json> plus the Llama-style switches of the same code base
(--swiglu --normalization rmsnorm --use-rotary-position-embeddings
--untie-embeddings-and-output-weights --disable-bias-linear --ffn-hidden-size
--num-key-value-heads --no-query-key-layer-scaling).  Unknown flags warn and are ignored.

Training loop, log line (loss, lr, grad-norm, ms/iter, samples/s, tokens/s, TFLOPs), eval
every `--eval-interval`, checkpoints in the Megatron-DeepSpeed layout every
`--save-interval`, auto-resume from `--load`.  `B2F_TRAIN_ITERS` caps --train-iters and
`B2F_OVERRIDE_ARGS` appends flags (the reference's 345M x 500 000 iterations are not a test).
"""
import argparse
import json
import math
import os
import sys
import time

import torch
import torch.distributed as dist

from b200flow.data.gpt_dataset import build_train_valid_test_blended, synthetic_corpus
from b200flow.data.indexed_dataset import MMapIndexedDataset
from b200flow.engines.ds_engine import initialize
from b200flow.models.gpt import GPTConfig
from b200flow.obs.tracking import Tracker, maybe_inject_fault
from b200flow.parallel.groups import ParallelState, init_distributed


def parse_args():
    p = argparse.ArgumentParser(allow_abbrev=False)
    a = p.add_argument
    a("--num-layers", type=int, default=24); a("--hidden-size", type=int, default=1024)
    a("--num-attention-heads", type=int, default=16); a("--num-key-value-heads", type=int, default=None)
    a("--ffn-hidden-size", type=int, default=None)
    a("--seq-length", type=int, default=1024); a("--max-position-embeddings", type=int, default=1024)
    a("--micro-batch-size", type=int, default=4); a("--global-batch-size", type=int, default=None)
    a("--lr", type=float, default=1.5e-4); a("--min-lr", type=float, default=1e-5)
    a("--train-iters", type=int, default=500000); a("--lr-decay-iters", type=int, default=None)
    a("--lr-decay-style", default="cosine"); a("--lr-warmup-fraction", type=float, default=0.01)
    a("--lr-warmup-iters", type=int, default=None)
    a("--weight-decay", type=float, default=0.01); a("--clip-grad", type=float, default=1.0)
    a("--adam-beta1", type=float, default=0.9); a("--adam-beta2", type=float, default=0.999)
    a("--adam-eps", type=float, default=1e-8)
    a("--fp16", action="store_true"); a("--bf16", action="store_true")
    a("--hidden-dropout", type=float, default=0.1); a("--attention-dropout", type=float, default=0.1)
    a("--init-method-std", type=float, default=0.02); a("--layernorm-epsilon", type=float, default=1e-5)
    a("--exit-interval", type=int, default=None)
    a("--tensor-model-parallel-size", type=int, default=1)
    a("--pipeline-model-parallel-size", type=int, default=1)
    a("--sequence-parallel", action="store_true"); a("--ds-sequence-parallel-size", type=int, default=1)
    a("--swiglu", action="store_true"); a("--normalization", default="layernorm")
    a("--use-rotary-position-embeddings", action="store_true")
    a("--untie-embeddings-and-output-weights", action="store_true")
    a("--disable-bias-linear", action="store_true")
    a("--no-query-key-layer-scaling", action="store_true")
    a("--use-flash-attn-v2", action="store_true"); a("--checkpoint-activations", action="store_true")
    a("--vocab-size", type=int, default=None); a("--make-vocab-size-divisible-by", type=int, default=128)
    a("--data-path", nargs="*", default=None); a("--vocab-file", default=None); a("--merge-file", default=None)
    a("--data-impl", default="mmap"); a("--split", default="949,50,1"); a("--data-cache-path", default=None)
    a("--tokenizer-type", default="GPT2BPETokenizer")
    a("--log-interval", type=int, default=100); a("--save-interval", type=int, default=None)
    a("--eval-interval", type=int, default=1000); a("--eval-iters", type=int, default=10)
    a("--distributed-backend", default="nccl"); a("--save", default=None); a("--load", default=None)
    a("--tensorboard-dir", default=None); a("--seed", type=int, default=1234)
    a("--no-save-optim", action="store_true"); a("--no-save-rng", action="store_true")
    a("--no-load-optim", action="store_true"); a("--no-load-rng", action="store_true"); a("--finetune", action="store_true")
    a("--deepspeed", action="store_true"); a("--deepspeed_config", default=None)
    a("--zero-stage", type=int, default=None); a("--no-pipeline-parallel", action="store_true")
    a("--local_rank", "--local-rank", type=int, default=None)
    # DeepSpeed-MoE flags of the pinned commit (examples_deepspeed/MoE/ds_pretrain_gpt_*_MoE*.sh upstream)
    a("--num-experts", type=int, nargs="+", default=[0]); a("--topk", type=int, default=1)
    a("--expert-interval", type=int, default=2); a("--moe-expert-parallel-size", type=int, default=1)
    a("--moe-loss-coeff", type=float, default=0.01); a("--moe-train-capacity-factor", type=float, default=1.0)
    a("--moe-eval-capacity-factor", type=float, default=1.0); a("--moe-min-capacity", type=int, default=4)
    a("--mlp-type", default="standard"); a("--create-moe-param-group", action="store_true")
    a("--disable-moe-token-dropping", action="store_true")
    import shlex
    argv = sys.argv[1:] + shlex.split(os.environ.get("B2F_OVERRIDE_ARGS", ""))  # test-size overrides
    args, unknown = p.parse_known_args(argv)
    if unknown and int(os.environ.get("RANK", 0)) == 0:
        print(f"WARNING: unknown arguments ignored: {unknown}", flush=True)
    return args


def padded_vocab(n, multiple, tp):
    m = multiple * tp
    return (n + m - 1) // m * m


def main():
    args = parse_args()
    backend = args.distributed_backend if torch.cuda.is_available() else "gloo"
    rank, local_rank, world = init_distributed(backend)
    tp, pp = args.tensor_model_parallel_size, args.pipeline_model_parallel_size
    n_experts = max(args.num_experts)
    ep = min(args.moe_expert_parallel_size, world // (tp * pp)) if n_experts > 1 else 1
    usp = args.ds_sequence_parallel_size        # DeepSpeed-Ulysses: ranks of a group share the micro-batches
    ps = ParallelState.initialize(tp=tp, pp=pp, ep=ep, sp=usp) if world > 1 else None
    dp = world // (tp * pp * usp)                # replicas that read different samples
    ds = json.load(open(args.deepspeed_config)) if args.deepspeed_config else {}
    micro = int(ds.get("train_micro_batch_size_per_gpu", args.micro_batch_size))
    gas = int(ds.get("gradient_accumulation_steps", 1))
    if args.global_batch_size:
        gas = max(1, args.global_batch_size // (micro * dp))
    train_iters = min(args.train_iters, int(os.environ.get("B2F_TRAIN_ITERS", args.train_iters)))

    vocab = args.vocab_size
    if vocab is None and args.vocab_file and os.path.exists(args.vocab_file):
        vocab = len(json.load(open(args.vocab_file)))
    vocab = padded_vocab(vocab or 50257, args.make_vocab_size_divisible_by, tp)
    llama = args.swiglu or args.normalization.lower() == "rmsnorm"
    cfg = GPTConfig(
        vocab_size=vocab, hidden=args.hidden_size, layers=args.num_layers, heads=args.num_attention_heads,
        kv_heads=args.num_key_value_heads, ffn=args.ffn_hidden_size or 4 * args.hidden_size,
        max_seq=args.max_position_embeddings, norm="rms" if args.normalization.lower() == "rmsnorm" else "ln",
        act="swiglu" if args.swiglu else "gelu", rope=args.use_rotary_position_embeddings,
        learned_pos=not args.use_rotary_position_embeddings, bias=not args.disable_bias_linear,
        tie_embeddings=not args.untie_embeddings_and_output_weights, name="llama" if llama else "gpt2",
        dropout=args.hidden_dropout, attention_dropout=args.attention_dropout, init_std=args.init_method_std,
        eps=args.layernorm_epsilon)
    if n_experts > 1:
        cfg.num_experts, cfg.moe_every, cfg.moe_top_k = n_experts, args.expert_interval, args.topk
        cfg.aux_loss_coeff, cfg.moe_min_capacity = args.moe_loss_coeff, args.moe_min_capacity
        # without token dropping every token must fit even if all of them pick one expert
        cfg.capacity_factor = float(n_experts) if args.disable_moe_token_dropping else args.moe_train_capacity_factor
        cfg.moe_eval_capacity_factor = args.moe_eval_capacity_factor
        cfg.name = "gpt-moe"
    decay_iters = args.lr_decay_iters or args.train_iters
    warm = args.lr_warmup_iters if args.lr_warmup_iters is not None else int(args.lr_warmup_fraction * decay_iters)
    ds_cfg = {
        "train_micro_batch_size_per_gpu": micro, "gradient_accumulation_steps": gas,
        "zero_optimization": ds.get("zero_optimization", {"stage": 1}),
        "gradient_clipping": args.clip_grad,
        "optimizer": {"params": {"lr": args.lr, "betas": (args.adam_beta1, args.adam_beta2),
                                 "eps": args.adam_eps, "weight_decay": args.weight_decay}},
        "scheduler": {"params": {"max_lr": args.lr, "min_lr": args.min_lr, "warmup_steps": warm,
                                 "decay_steps": decay_iters, "decay_style": args.lr_decay_style}},
        # Megatron's --sequence-parallel only changes where activations live between the TP linears; the
        # numerics are the same.  The engine therefore shards the sequence whenever TP > 1 (that is the form
        # the fused all-gather->GEMM / GEMM->reduce-scatter kernels implement) unless B2F_AUTO_SP=0 asks for
        # the replicated-activation + all-reduce form of a run without the flag.
        "sequence_parallel": bool(args.sequence_parallel) or tp == 1 or os.environ.get("B2F_AUTO_SP", "1") != "0",
        "activation_checkpointing": {"enabled": args.checkpoint_activations},
        "fused_collectives": os.environ.get("B2F_FUSED", "1") == "1",
    }
    if ds.get("fp16", {}).get("enabled") or (args.fp16 and not ds.get("bf16", {}).get("enabled")):
        ds_cfg["fp16"] = {**ds.get("fp16", {}), "enabled": True}    # loss-scale state machine (skipped steps)
    engine = initialize(cfg, ds_cfg, ps, seed=args.seed)
    engine.args = vars(args)
    dev = engine.device

    # ---- data: TP rank 0 semantics collapse to "every rank of a DP replica reads the same sample"
    prefix = args.data_path[-1] if args.data_path else None
    # rank 0 decides whether a corpus has to be generated and everyone follows its decision: ranks looking at the
    # directory on their own could see rank 0's half-written files, take them for a corpus and skip the barrier
    need = [bool(not prefix or not MMapIndexedDataset.exists(prefix))]
    if world > 1:
        dist.broadcast_object_list(need, src=0)
    if need[0]:
        home = os.environ.get("HOME", "/tmp")
        prefix = os.path.join(home, "synthetic-corpus", "gpt2_text_document")
        if rank == 0 and not MMapIndexedDataset.exists(prefix):
            print(f"[b200flow] no corpus at --data-path; generating a synthetic one at {prefix}", flush=True)
            tmp = prefix + f".tmp{os.getpid()}"
            synthetic_corpus(tmp, min(vocab, 50257), num_docs=1500, mean_len=max(256, args.seq_length // 2))
            os.replace(tmp + ".bin", prefix + ".bin")      # complete files appear under the final names
            os.replace(tmp + ".idx", prefix + ".idx")
        if world > 1:
            dist.barrier()
    gbs = micro * gas * dp
    n_eval = (train_iters // max(1, args.eval_interval) + 1) * args.eval_iters * gbs
    data_path = [prefix]
    if args.data_path and len(args.data_path) > 1:      # weight / prefix pairs: a blend of corpora
        from b200flow.data.gpt_dataset import parse_data_path

        if all(MMapIndexedDataset.exists(pfx) for pfx in parse_data_path(args.data_path)[1]):
            data_path = list(args.data_path)
    train_ds, valid_ds, _ = build_train_valid_test_blended(data_path, args.split, args.seq_length,
                                                           (train_iters * gbs, max(n_eval, gbs), 0), args.seed,
                                                           args.data_cache_path)
    dp_rank = (ps.dp_rank // usp) if ps else 0

    def batches(dataset, start_step):
        step = start_step
        while True:
            for g in range(gas):
                idx0 = (step * gas + g) * micro * dp + dp_rank * micro
                toks = torch.stack([dataset[idx0 + i]["text"] for i in range(micro)])
                toks = toks.clamp_(max=vocab - 1).to(dev, non_blocking=True)
                yield toks[:, :-1].contiguous(), toks[:, 1:].contiguous()
            step += 1

    start = 0
    if not args.load and args.save and os.environ.get("B2F_AUTO_RESUME") == "1":
        args.load = args.save          # launcher --auto-resume: continue from what this job saved before
    if args.load and os.path.exists(os.path.join(args.load, "latest")):
        path, client = engine.load_checkpoint(args.load, load_optimizer_states=not args.no_load_optim,
                                              load_rng=not args.no_load_rng, finetune=args.finetune)
        start = engine.global_steps
        if rank == 0:
            print(f" successfully loaded checkpoint from {path} at iteration {start}", flush=True)
    tracker = Tracker(args.tensorboard_dir, "megatron") if (args.tensorboard_dir and rank == 0) else None
    it = batches(train_ds, start)
    flops_tok = cfg.flops_per_token(args.seq_length)
    t_last, loss_acc, n_acc = time.time(), 0.0, 0
    for iteration in range(start + 1, train_iters + 1):
        maybe_inject_fault(iteration)
        loss = engine.train_batch(it)
        loss_acc += float(loss)
        n_acc += 1
        if iteration % args.log_interval == 0 or iteration == train_iters:
            torch.cuda.synchronize() if dev.type == "cuda" else None
            dt = (time.time() - t_last) / n_acc
            tok_s = gbs * args.seq_length / dt
            line = (f" iteration {iteration:8d}/{train_iters:8d} | consumed samples: {iteration * gbs:12d} | "
                    f"elapsed time per iteration (ms): {dt * 1e3:.1f} | learning rate: {engine.optimizer.lr:.3E} | "
                    f"global batch size: {gbs:5d} | lm loss: {loss_acc / n_acc:.6E} | "
                    + (f"loss scale: {engine.loss_scale:.1f} | " if engine.loss_scaler is not None else "")
                    + f"grad norm: {engine.grad_norm():.3f} | "
                    + (f"number of skipped iterations: {engine.skipped_steps:3d} | "
                       if engine.loss_scaler is not None else "") +
                    f"samples per second: {gbs / dt:.3f} | tokens per second: {tok_s:.1f} | "
                    f"TFLOPs: {flops_tok * tok_s / world / 1e12:.2f} |")
            if rank == 0:
                print(line, flush=True)
                if tracker:
                    tracker.log({"lm loss": loss_acc / n_acc, "learning-rate": engine.optimizer.lr,
                                 "grad-norm": engine.grad_norm(), "tokens-per-second": tok_s}, iteration)
            t_last, loss_acc, n_acc = time.time(), 0.0, 0
        if valid_ds is not None and args.eval_interval and iteration % args.eval_interval == 0:
            vit = batches(valid_ds, iteration // args.eval_interval)
            with torch.no_grad():
                vl = sum(float(engine.eval_batch(vit)) for _ in range(args.eval_iters)) / args.eval_iters
            if rank == 0:
                print(f" validation loss at iteration {iteration} | lm loss value: {vl:.6E} | "
                      f"lm loss PPL: {math.exp(min(20, vl)):.6E} |", flush=True)
        if args.save and args.save_interval and iteration % args.save_interval == 0:
            engine.save_checkpoint(args.save, save_optimizer=not args.no_save_optim, save_rng=not args.no_save_rng)
            if rank == 0:
                print(f"  successfully saved checkpoint at iteration {iteration:7d} to {args.save}", flush=True)
        if args.exit_interval and iteration % args.exit_interval == 0:
            if rank == 0:
                print(f" exiting program at iteration {iteration}", flush=True)
            train_iters = iteration
            break
    if args.save and (not args.save_interval or train_iters % args.save_interval != 0):
        engine.save_checkpoint(args.save, save_optimizer=not args.no_save_optim, save_rng=not args.no_save_rng)
    if tracker:
        tracker.finish()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
