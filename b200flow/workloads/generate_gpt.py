"""Text generation from a Megatron-DeepSpeed-layout checkpoint (Megatron's `tools/generate_samples_gpt.py`).

    python -m b200flow.workloads.generate_gpt --load <dir> --prompt "The launcher" --max-new-tokens 40 \
        [--vocab-file gpt2-vocab.json --merge-file gpt2-merges.txt] [--temperature 0.8 --top-k 40]

Single process, tensor-parallel size 1 (convert with `python -m b200flow.ckpt.reshard --target-tp 1` first).  The model
geometry comes from the arguments stored in the checkpoint by `pretrain_gpt.py`.  Without tokenizer files the prompt is
a list of token ids and so is the output.
"""
from __future__ import annotations

import argparse

import torch

from b200flow.engines.ds_engine import initialize
from b200flow.models.gpt import GPTConfig


def config_from_args(a: dict) -> GPTConfig:
    llama = bool(a.get("swiglu")) or str(a.get("normalization", "")).lower() == "rmsnorm"
    vocab = a.get("padded_vocab_size") or a.get("vocab_size") or 50304
    return GPTConfig(vocab_size=int(vocab), hidden=a["hidden_size"], layers=a["num_layers"], heads=a["num_attention_heads"],
                     kv_heads=a.get("num_key_value_heads"), ffn=a.get("ffn_hidden_size") or 4 * a["hidden_size"],
                     max_seq=a["max_position_embeddings"], norm="rms" if str(a.get("normalization", "")).lower() == "rmsnorm" else "ln",
                     act="swiglu" if a.get("swiglu") else "gelu", rope=bool(a.get("use_rotary_position_embeddings")),
                     learned_pos=not a.get("use_rotary_position_embeddings"), bias=not a.get("disable_bias_linear"),
                     tie_embeddings=not a.get("untie_embeddings_and_output_weights"), name="llama" if llama else "gpt2",
                     eps=a.get("layernorm_epsilon", 1e-5))


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--load", required=True)
    ap.add_argument("--prompt", default="")
    ap.add_argument("--prompt-ids", type=int, nargs="*", default=None)
    ap.add_argument("--max-new-tokens", type=int, default=32)
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--top-k", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--vocab-file", default=None)
    ap.add_argument("--merge-file", default=None)
    args = ap.parse_args(argv)
    from pathlib import Path

    root = Path(args.load)
    tag = (root / "latest").read_text().strip()
    client = torch.load(root / tag / "mp_rank_00_model_states.pt", map_location="cpu", weights_only=False)
    margs = client.get("args")
    assert margs, "the checkpoint carries no model arguments (it was not written by pretrain_gpt.py)"
    if margs.get("tensor_model_parallel_size", 1) != 1 or margs.get("pipeline_model_parallel_size", 1) != 1:
        raise SystemExit("generation runs at TP = PP = 1: reshard the checkpoint first (python -m b200flow.ckpt.reshard)")
    cfg = config_from_args(dict(margs, padded_vocab_size=next(
        (v.shape[0] for k, v in (client.get("module") or {}).items() if k == "embed.wte"), None)))
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 1}, device=dev)
    eng.load_checkpoint(args.load, load_optimizer_states=False)
    tok = None
    if args.vocab_file and args.merge_file:
        from b200flow.data.tokenizer import GPT2BPETokenizer

        tok = GPT2BPETokenizer(args.vocab_file, args.merge_file)
    ids = args.prompt_ids if args.prompt_ids is not None else (tok.tokenize(args.prompt) if tok else [0])
    g = torch.Generator(device="cpu").manual_seed(args.seed)
    out = eng.model.generate(torch.tensor([ids], device=dev), args.max_new_tokens, args.temperature, args.top_k,
                             eos_id=tok.eod if tok else None, generator=g if dev.type == "cpu" else None)
    new = out[0].tolist()
    print("token ids:", new)
    if tok:
        print(tok.detokenize(new))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
