"""GPT-2 / Llama decoder-only transformer on the b200flow sm_100a ops.

One module covers both families the reference trains through Megatron-DeepSpeed
(`examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:39-53` for GPT-2
345M; Llama-2 7B is BASELINE.json config 4): LayerNorm+GELU+learned positions+tied
embeddings for GPT-2, RMSNorm+SwiGLU+RoPE+untied head for Llama.

Activations are ``[tokens, hidden]`` with tokens in sequence-major ``[s, b]`` order
(Megatron's layout) so a sequence-parallel shard is a contiguous row range.  Every
linear goes through a :class:`~b200flow.parallel.tensor_parallel.TPContext`, which
is the identity on one GPU and the fused all-gather→GEMM / GEMM→reduce-scatter
kernels under tensor + sequence parallelism.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
from torch import nn

from ..ops import attention as A
from ..ops import elementwise as E
from ..ops import gemm as G
from ..ops import norm as N


@dataclass
class GPTConfig:
    vocab_size: int = 32000
    hidden: int = 4096
    layers: int = 32
    heads: int = 32
    kv_heads: int | None = None
    ffn: int = 11008
    max_seq: int = 4096
    norm: str = "rms"            # "rms" | "ln"
    act: str = "swiglu"          # "swiglu" | "gelu"
    rope: bool = True
    rope_base: float = 10000.0
    learned_pos: bool = False
    bias: bool = False
    tie_embeddings: bool = False
    eps: float = 1e-5
    init_std: float = 0.02
    dropout: float = 0.0          # --hidden-dropout (embeddings, attention output, MLP output)
    attention_dropout: float = 0.0   # --attention-dropout (softmax probabilities, inside the flash kernels)
    name: str = "llama"
    num_experts: int = 0          # > 0: MoE FFN every `moe_every`-th layer (DeepSpeed --num-experts)
    moe_every: int = 2            # --expert-interval
    capacity_factor: float = 1.0
    aux_loss_coeff: float = 0.01
    moe_top_k: int = 1            # --topk (DeepSpeed TopKGate: 1 or 2)
    moe_min_capacity: int = 4     # --moe-min-capacity
    moe_eval_capacity_factor: float | None = None   # --moe-eval-capacity-factor (None: same as training)
    moe_fp8: bool = False         # expert GEMMs (forward + dgrad) in e4m3

    def __post_init__(self):
        if self.kv_heads is None:
            self.kv_heads = self.heads

    @property
    def head_dim(self) -> int:
        return self.hidden // self.heads

    def param_count(self) -> int:
        d = self.head_dim
        qkv = self.hidden * (self.heads + 2 * self.kv_heads) * d
        proj = self.hidden * self.hidden
        mlp = self.hidden * self.ffn * (3 if self.act == "swiglu" else 2)
        per_layer = qkv + proj + mlp + 2 * self.hidden
        emb = self.vocab_size * self.hidden * (1 if self.tie_embeddings else 2)
        total = self.layers * per_layer + emb + self.hidden
        if self.num_experts:
            total += (self.layers // self.moe_every) * (self.num_experts - 1) * mlp
        return total

    def flops_per_token(self, seq: int) -> float:
        """Model FLOPs (fwd+bwd) per token, causal attention counted at half."""
        d = self.head_dim
        lin = 2 * (self.hidden * (self.heads + 2 * self.kv_heads) * d + self.hidden * self.hidden
                   + self.hidden * self.ffn * (3 if self.act == "swiglu" else 2))
        attn = 2 * 2 * seq * self.hidden * 0.5
        head = 2 * self.hidden * self.vocab_size
        return 3.0 * (self.layers * (lin + attn) + head)


PRESETS = {
    # Llama-2 7B (BASELINE.json config 4)
    "llama2-7b": dict(vocab_size=32000, hidden=4096, layers=32, heads=32, ffn=11008, max_seq=4096),
    # GPT-2 345M exactly as examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:39-45
    "gpt2-345m": dict(vocab_size=50304, hidden=1024, layers=24, heads=16, ffn=4096, max_seq=1024,
                      norm="ln", act="gelu", rope=False, learned_pos=True, bias=True,
                      tie_embeddings=True, name="gpt2"),
    # GPT-MoE 1.3B, 8 experts (BASELINE.json config 5; [UP] ds_pretrain_gpt_1.3B_MoE*.sh)
    "gpt-moe-1.3b": dict(vocab_size=50304, hidden=2048, layers=24, heads=16, ffn=8192, max_seq=2048,
                         norm="ln", act="gelu", rope=False, learned_pos=True, bias=True,
                         tie_embeddings=True, name="gpt-moe", num_experts=8, moe_every=2),
    "gpt-moe-tiny": dict(vocab_size=1024, hidden=128, layers=4, heads=4, ffn=256, max_seq=128,
                         norm="ln", act="gelu", rope=False, learned_pos=True, bias=True,
                         tie_embeddings=True, name="gpt-moe", num_experts=4, moe_every=2),
    "llama-tiny": dict(vocab_size=1024, hidden=256, layers=2, heads=2, ffn=512, max_seq=256),
    "gpt2-tiny": dict(vocab_size=1024, hidden=256, layers=2, heads=4, ffn=1024, max_seq=256,
                      norm="ln", act="gelu", rope=False, learned_pos=True, bias=True,
                      tie_embeddings=True, name="gpt2"),
}


def config_from_preset(name: str, **overrides) -> GPTConfig:
    kw = dict(PRESETS[name])
    kw.update(overrides)
    return GPTConfig(**kw)


class _QKVAttention(torch.autograd.Function):
    """RoPE (in place on the fused QKV buffer) + flash attention, one autograd node.

    qkv: [s, b, heads + 2*kv_heads, d].  Returns o: [s*b, heads*d] (sequence-major)."""

    @staticmethod
    def forward(ctx, qkv, heads, kv_heads, causal, scale, rope_tables, kv_lens, dropout_p=0.0, seed=0):
        s, b, nh, d = qkv.shape
        if rope_tables is not None:
            x = qkv.view(s * b, nh, d)[:, : heads + kv_heads]
            E.rope_inplace_raw(x, rope_tables[0], rope_tables[1], s, False, 0, pos_div=b)
        v = qkv.permute(1, 0, 2, 3)  # [b, s, nh, d] view
        q, k, vv = v[:, :, :heads], v[:, :, heads:heads + kv_heads], v[:, :, heads + kv_heads:]
        o_sb = torch.empty((s, b, heads, d), dtype=qkv.dtype, device=qkv.device)
        o = o_sb.permute(1, 0, 2, 3)
        lse = torch.empty((b, heads, s), dtype=torch.float32, device=qkv.device)
        args = [A.ptr(q), A.ptr(k), A.ptr(vv), A.ptr(o), A.ptr(lse), A.c_int(b),
                A.c_int(s), A.c_int(heads), A.c_int(kv_heads), A.c_int(d),
                A.c_ll(q.stride(1)), A.c_ll(q.stride(0)), A.c_ll(k.stride(1)), A.c_ll(k.stride(0)),
                A.c_ll(vv.stride(1)), A.c_ll(vv.stride(0)), A.c_ll(o.stride(1)), A.c_ll(o.stride(0)),
                A.c_float(scale), A.c_int(int(causal)), A.ptr(kv_lens)]
        if dropout_p > 0:
            A._lib.call("b2f_attn_fwd_drop", *args, A.c_float(dropout_p), A.ctypes.c_uint(seed & 0xFFFFFFFF),
                        A.stream_ptr())
        else:
            A._lib.call("b2f_attn_fwd", *args, A.stream_ptr())
        ctx.save_for_backward(qkv, o_sb, lse)
        ctx.cfg = (heads, kv_heads, causal, scale, rope_tables, kv_lens)
        ctx.drop = (dropout_p, seed)
        return o_sb.view(s * b, heads * d)

    @staticmethod
    def backward(ctx, do):
        qkv, o_sb, lse = ctx.saved_tensors
        heads, kv_heads, causal, scale, rope_tables, kv_lens = ctx.cfg
        s, b, nh, d = qkv.shape
        do_sb = do.reshape(s, b, heads, d)
        if not do_sb.is_contiguous():
            do_sb = do_sb.contiguous()
        dqkv = torch.empty_like(qkv)
        v = qkv.permute(1, 0, 2, 3)
        g = dqkv.permute(1, 0, 2, 3)
        A.attn_bwd_raw(v[:, :, :heads], v[:, :, heads:heads + kv_heads], v[:, :, heads + kv_heads:],
                       o_sb.permute(1, 0, 2, 3), do_sb.permute(1, 0, 2, 3), lse, causal, scale, kv_lens,
                       dq=g[:, :, :heads], dk=g[:, :, heads:heads + kv_heads],
                       dv=g[:, :, heads + kv_heads:], dropout_p=ctx.drop[0], seed=ctx.drop[1])
        if rope_tables is not None:
            x = dqkv.view(s * b, nh, d)[:, : heads + kv_heads]
            E.rope_inplace_raw(x, rope_tables[0], rope_tables[1], s, True, 0, pos_div=b)
        return dqkv, None, None, None, None, None, None, None, None


def qkv_attention(qkv, heads, kv_heads, causal, scale, rope_tables=None, kv_lens=None, dropout_p=0.0, seed=0):
    if not qkv.is_cuda:  # CPU oracle path (gloo plumbing tests)
        from ..ops import ref
        s, b, nh, d = qkv.shape
        x = qkv
        if rope_tables is not None:
            pos = torch.arange(s * b, device=qkv.device) // b
            rot = ref.rope(qkv.reshape(s * b, nh, d)[:, : heads + kv_heads], rope_tables[0],
                           rope_tables[1], pos)
            x = torch.cat([rot, qkv.reshape(s * b, nh, d)[:, heads + kv_heads:]], 1).view(s, b, nh, d)
        v = x.permute(1, 0, 2, 3)
        o = ref.attention(v[:, :, :heads], v[:, :, heads:heads + kv_heads], v[:, :, heads + kv_heads:],
                          causal, scale, kv_lens, dropout_p, seed)
        return o.permute(1, 0, 2, 3).reshape(s * b, heads * d)
    return _QKVAttention.apply(qkv, heads, kv_heads, causal, scale, rope_tables, kv_lens, float(dropout_p), int(seed))


class TransformerLayer(nn.Module):
    """Pre-norm decoder block. `tp` supplies the (possibly fused-collective) linears."""

    def __init__(self, cfg: GPTConfig, tp, layer_idx: int, dtype=torch.bfloat16, device=None, ep=None,
                 fused_moe: bool = True):
        super().__init__()
        self.cfg, self.tp, self.layer_idx = cfg, tp, layer_idx
        self.ulysses_group = None      # set by the engine (`sequence_parallel_size` > 1)
        self.is_moe = cfg.num_experts > 0 and (layer_idx + 1) % cfg.moe_every == 0
        d = cfg.head_dim
        assert cfg.heads % tp.world == 0 and cfg.kv_heads % tp.world == 0 and cfg.ffn % tp.world == 0
        self.h_local = cfg.heads // tp.world
        self.kv_local = cfg.kv_heads // tp.world
        ffn_local = cfg.ffn // tp.world
        kw = dict(dtype=dtype, device=device)
        H = cfg.hidden
        self.ln1_w = nn.Parameter(torch.ones(H, **kw))
        self.ln2_w = nn.Parameter(torch.ones(H, **kw))
        if cfg.norm == "ln":
            self.ln1_b = nn.Parameter(torch.zeros(H, **kw))
            self.ln2_b = nn.Parameter(torch.zeros(H, **kw))
        else:
            self.ln1_b = self.ln2_b = None
        std = cfg.init_std
        out_std = std / math.sqrt(2.0 * cfg.layers)
        self.wqkv = nn.Parameter(torch.empty((self.h_local + 2 * self.kv_local) * d, H, **kw).normal_(0, std))
        self.wo = nn.Parameter(torch.empty(H, self.h_local * d, **kw).normal_(0, out_std))
        n_up = 2 * ffn_local if cfg.act == "swiglu" else ffn_local
        self.moe = None
        if self.is_moe:
            from ..parallel.moe import MoELayer

            assert tp.world == 1 and cfg.act == "gelu", "MoE layers: TP=1, GELU experts (GPT-MoE recipe)"
            self.moe = MoELayer(H, cfg.ffn, cfg.num_experts, ep, cfg.capacity_factor, act="gelu",
                                aux_loss_coeff=cfg.aux_loss_coeff, init_std=std, dtype=dtype, device=device,
                                fused=fused_moe, fp8=cfg.moe_fp8, top_k=cfg.moe_top_k,
                                min_capacity=cfg.moe_min_capacity, eval_capacity_factor=cfg.moe_eval_capacity_factor)
            self.w1 = self.w2 = None
        else:
            self.w1 = nn.Parameter(torch.empty(n_up, H, **kw).normal_(0, std))       # gate|up or fc1
            self.w2 = nn.Parameter(torch.empty(H, ffn_local, **kw).normal_(0, out_std))
        if cfg.bias:
            self.bqkv = nn.Parameter(torch.zeros(self.wqkv.shape[0], **kw))
            self.bo = nn.Parameter(torch.zeros(H, **kw))
            self.b1 = None if self.is_moe else nn.Parameter(torch.zeros(n_up, **kw))
            self.b2 = None if self.is_moe else nn.Parameter(torch.zeros(H, **kw))
        else:
            self.bqkv = self.bo = self.b1 = self.b2 = None
        # parameters replicated across the TP group need their grads all-reduced there
        for p in (self.ln1_w, self.ln2_w, self.ln1_b, self.ln2_b, self.bo, self.b2):
            if p is not None:
                p.tp_replicated = True

    def _dropout_add(self, x, residual, p, seed):
        """residual + dropout(x).  Sequence-parallel activations are different rows on every TP rank and take a
        per-rank mask; replicated activations must get the same mask everywhere (Megatron's two RNG streams)."""
        if self.tp.world > 1 and self.tp.sequence_parallel:
            seed = seed + 7919 * (self.tp.rank + 1)
        return E.bias_dropout_add(x, None, residual, p, seed)

    def forward(self, h, seq: int, batch: int, rope_tables=None, kv_lens=None, drop_seed: int = 0):
        cfg, tp = self.cfg, self.tp
        rms = cfg.norm == "rms"
        d = cfg.head_dim
        p = cfg.dropout if self.training else 0.0
        seed = drop_seed + 2 * self.layer_idx
        # --attention-dropout: same mask for every TP rank's own heads (the head index inside the kernel is local, so the
        # rank goes into the seed); a recomputed forward draws it again from the per-micro-batch seed
        pa = cfg.attention_dropout if self.training else 0.0
        aseed = (drop_seed * 2654435761 + 40503 * self.layer_idx + 7919 * tp.rank + 1) & 0x7FFFFFFF
        y = N.norm(h, self.ln1_w, self.ln1_b, eps=cfg.eps, rms=rms)
        qkv = tp.column_linear(y, self.wqkv, self.bqkv)              # [s*b, (h+2kv)*d] full sequence
        if self.ulysses_group is not None:     # this rank holds a sequence slice: regroup to whole sequences of fewer heads
            from ..parallel.ulysses import ulysses_qkv_attention

            o = ulysses_qkv_attention(qkv.view(seq, batch, self.h_local + 2 * self.kv_local, d), self.h_local,
                                      self.kv_local, True, 1.0 / math.sqrt(d), self.ulysses_group, qkv_attention,
                                      rope_tables, kv_lens)
        else:
            o = qkv_attention(qkv.view(seq, batch, self.h_local + 2 * self.kv_local, d), self.h_local,
                              self.kv_local, True, 1.0 / math.sqrt(d), rope_tables, kv_lens, pa, aseed)
        if p > 0:      # --hidden-dropout: the residual add moves out of the GEMM epilogue, behind the mask
            h = self._dropout_add(tp.row_linear(o, self.wo, self.bo), h, p, seed)
        else:
            h = tp.row_linear(o, self.wo, self.bo, residual=h)       # residual fused in the epilogue
        y = N.norm(h, self.ln2_w, self.ln2_b, eps=cfg.eps, rms=rms)
        if self.moe is not None:
            return h + self.moe(y)
        if cfg.act == "swiglu":
            a = E.swiglu(tp.column_linear(y, self.w1, self.b1))
        else:
            a = tp.column_linear_gelu(y, self.w1, self.b1)
        if p > 0:
            return self._dropout_add(tp.row_linear(a, self.w2, self.b2), h, p, seed + 1)
        h = tp.row_linear(a, self.w2, self.b2, residual=h)
        return h


class _ScaleGrad(torch.autograd.Function):
    """Identity forward; the gradient is multiplied by a constant."""

    @staticmethod
    def forward(ctx, x, scale):
        ctx.scale = scale
        return x.view_as(x)

    @staticmethod
    def backward(ctx, dy):
        return dy * ctx.scale, None


class Embedding(nn.Module):
    def __init__(self, cfg: GPTConfig, tp, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.cfg, self.tp = cfg, tp
        self.grad_scale = 1.0      # see GPTModel: tied table under tensor parallelism
        self.pos_offset = 0        # first position of this rank's sequence slice (Ulysses)
        kw = dict(dtype=dtype, device=device)
        self.wte = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden, **kw).normal_(0, cfg.init_std))
        if cfg.learned_pos:
            self.wpe = nn.Parameter(torch.empty(cfg.max_seq, cfg.hidden, **kw).normal_(0, cfg.init_std))
        else:
            self.wpe = None

    def forward(self, tokens):
        """tokens: [b, s] int64 -> [s*b, hidden] (sequence-major; SP shard taken by tp.scatter)."""
        b, s = tokens.shape
        x = torch.nn.functional.embedding(tokens.t().contiguous(), self.wte)  # [s, b, h]
        if self.grad_scale != 1.0:
            x = _ScaleGrad.apply(x, self.grad_scale)
        if self.wpe is not None:
            x = x + self.wpe[self.pos_offset:self.pos_offset + s].unsqueeze(1)
        return self.tp.scatter_sequence(x.view(s * b, self.cfg.hidden))


class LMHead(nn.Module):
    def __init__(self, cfg: GPTConfig, tp, tied_weight=None, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.cfg, self.tp = cfg, tp
        kw = dict(dtype=dtype, device=device)
        self.lnf_w = nn.Parameter(torch.ones(cfg.hidden, **kw))
        self.lnf_w.tp_replicated = True
        if cfg.norm == "ln":
            self.lnf_b = nn.Parameter(torch.zeros(cfg.hidden, **kw))
            self.lnf_b.tp_replicated = True
        else:
            self.lnf_b = None
        self.tied_slice = None       # (lo, hi): this TP rank's vocabulary rows of a tied, replicated table
        if tied_weight is not None:
            self.weight = tied_weight
            if tp.world > 1:
                assert cfg.vocab_size % tp.world == 0
                n = cfg.vocab_size // tp.world
                self.tied_slice = (tp.rank * n, (tp.rank + 1) * n)
        else:
            assert cfg.vocab_size % tp.world == 0
            self.weight = nn.Parameter(
                torch.empty(cfg.vocab_size // tp.world, cfg.hidden, **kw).normal_(0, cfg.init_std))

    def forward(self, h, labels):
        """h: [tokens(_shard), hidden]; labels: [b, s] -> mean cross-entropy (scalar)."""
        cfg = self.cfg
        y = N.norm(h, self.lnf_w, self.lnf_b, eps=cfg.eps, rms=cfg.norm == "rms")
        tgt = labels.t().contiguous().view(-1)  # sequence-major, matches the activations
        w = self.weight
        if self.tied_slice is not None:      # vocab-parallel head over this rank's rows of the shared table
            lo, hi = self.tied_slice
            w = self.weight[lo:hi]
            mg = getattr(self.weight, "main_grad", None)
            if mg is not None:
                w.main_grad = mg[lo:hi]      # the wgrad GEMM accumulates straight into those rows
        return self.tp.lm_head_loss(y, w, tgt)


class GPTModel(nn.Module):
    """Full model (or one pipeline stage of it: `layer_range`, `pre`, `post`)."""

    def __init__(self, cfg: GPTConfig, tp=None, layer_range: tuple[int, int] | None = None,
                 pre: bool = True, post: bool = True, dtype=torch.bfloat16, device=None, ep=None,
                 fused_moe: bool = True):
        super().__init__()
        from ..parallel.tensor_parallel import SINGLE

        self.cfg = cfg
        self.tp = tp if tp is not None else SINGLE
        lo, hi = layer_range if layer_range is not None else (0, cfg.layers)
        self.layer_range = (lo, hi)
        self.embed = Embedding(cfg, self.tp, dtype, device) if pre else None
        self.layers = nn.ModuleList(
            [TransformerLayer(cfg, self.tp, i, dtype, device, ep, fused_moe) for i in range(lo, hi)])
        if post:
            tied = self.embed.wte if (cfg.tie_embeddings and pre) else None
            self.head = LMHead(cfg, self.tp, tied, dtype, device)
            if tied is not None and self.tp.world > 1:
                # One table, two uses, TP ranks: the lookup's gradient is complete (and equal) on every rank, the
                # head's only covers the rank's own vocabulary rows.  Scaling the lookup's gradient by 1/TP makes
                # the plain sum over the TP group (engine: `tp_sum_grad`) equal lookup + head gradient.
                self.embed.grad_scale = 1.0 / self.tp.world
                self.embed.wte.tp_sum_grad = True
        else:
            self.head = None
        self.recompute = False
        self.seq_mult = 1        # Ulysses: rotary tables cover the whole sequence, `seq` is this rank's slice
        self.drop_seed = 0       # set per micro-batch by the engine, so a recomputed forward draws the same masks

    @torch.no_grad()
    def logits(self, tokens: torch.Tensor) -> torch.Tensor:
        """tokens [b, s] -> next-token logits [b, s, vocab] (single process, no pipeline stages): the forward pass
        without the loss, for evaluation by hand and for `generate`."""
        assert self.embed is not None and self.head is not None and self.tp.world == 1
        was = self.training
        self.eval()
        try:
            b, s = tokens.shape
            h = self.forward(tokens=tokens)                                   # [s*b, hidden], sequence-major
            y = N.norm(h, self.head.lnf_w, self.head.lnf_b, eps=self.cfg.eps, rms=self.cfg.norm == "rms")
            w = self.head.weight
            out = G.linear(y, w) if y.is_cuda else torch.nn.functional.linear(y, w)
            return out.view(s, b, -1).transpose(0, 1)
        finally:
            self.train(was)

    @torch.no_grad()
    def generate(self, tokens: torch.Tensor, max_new_tokens: int, temperature: float = 0.0, top_k: int = 0,
                 eos_id: int | None = None, generator: torch.Generator | None = None) -> torch.Tensor:
        """Autoregressive sampling (Megatron's text-generation utilities: greedy when temperature is 0, otherwise
        temperature / top-k sampling).  The whole prefix is re-encoded for every new token — no key/value cache —
        which is fine for checking a trained model, not for serving."""
        out = tokens
        for _ in range(max_new_tokens):
            ctx = out[:, -self.cfg.max_seq:]
            pad = (-ctx.shape[1]) % 8 if ctx.is_cuda else 0                 # the attention kernel wants multiples of 8
            if pad and ctx.shape[1] + pad <= self.cfg.max_seq:
                ctx = torch.nn.functional.pad(ctx, (0, pad))
            else:
                pad = 0
            last = self.logits(ctx)[:, ctx.shape[1] - 1 - pad].float()        # causal: padding behind it is not seen
            if temperature <= 0:
                nxt = last.argmax(-1, keepdim=True)
            else:
                last = last / temperature
                if top_k > 0:
                    kth = last.topk(min(top_k, last.shape[-1]), dim=-1).values[:, -1:]
                    last = last.masked_fill(last < kth, float("-inf"))
                nxt = torch.multinomial(last.softmax(-1), 1, generator=generator)
            out = torch.cat([out, nxt], 1)
            if eos_id is not None and bool((nxt == eos_id).all()):
                break
        return out

    def _rope(self, seq, device):
        if not self.cfg.rope:
            return None
        return E.rope_tables(seq, self.cfg.head_dim, self.cfg.rope_base, device)

    def forward(self, tokens=None, labels=None, hidden=None, seq: int | None = None,
                batch: int | None = None, kv_lens=None):
        """First stage takes `tokens`; later stages take `hidden`. Last stage returns the loss."""
        if self.embed is not None:
            batch, seq = tokens.shape
            h = self.embed(tokens)
        else:
            h = hidden
        rope = self._rope(seq * self.seq_mult, h.device)
        drop = self.cfg.dropout if self.training else 0.0
        seed = int(self.drop_seed)
        if drop > 0 and self.embed is not None:      # Megatron applies the hidden dropout to the embeddings too
            es = seed - 1 + (7919 * (self.tp.rank + 1) if self.tp.world > 1 and self.tp.sequence_parallel else 0)
            h = E.bias_dropout_add(h, None, None, drop, es)
        for layer in self.layers:
            if self.recompute and torch.is_grad_enabled():
                h = torch.utils.checkpoint.checkpoint(layer, h, seq, batch, rope, kv_lens, seed,
                                                      use_reentrant=False)
            else:
                h = layer(h, seq, batch, rope, kv_lens, seed)
        if self.head is not None and labels is not None:
            loss = self.head(h, labels)
            if self.cfg.num_experts:
                aux = [l.moe.last_aux for l in self.layers if l.moe is not None and l.moe.last_aux is not None]
                if aux:
                    self.last_lm_loss = loss.detach()
                    loss = loss + self.cfg.aux_loss_coeff * torch.stack(aux).sum()
            return loss
        return h
