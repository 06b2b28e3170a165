"""Change the tensor-parallel size of a Megatron-DeepSpeed-layout checkpoint (weights only).

    python -m b200flow.ckpt.reshard --load <dir> --save <dir> --target-tp N [--tag global_stepK]

Upstream users do this with Megatron's checkpoint conversion tools when a run moves to a different GPU count;
the result is loaded with `--no-load-optim` (the ZeRO optimizer partitions belong to the old layout — the
data-parallel size alone can change without this tool, see `megatron_ds_layout._load_optimizer_state`).

Both file families of the layout are handled: `mp_rank_<TP>_model_states.pt` (whole module, no pipeline
parallelism) and `layer_<L>-model_<TP>-model_states.pt` (one file per layer under pipeline parallelism).  How a
tensor is split follows from its leaf name (`models/gpt.py`):

    wqkv, bqkv   rows grouped as [q heads | k heads | v heads]; each group is cut by heads
    w1, b1       rows; for SwiGLU the gate and up halves are cut separately
    wo, w2       columns (row-parallel)
    head weight  vocabulary rows — unless every source shard is identical (tied / replicated), then it stays whole
    everything else (norms, embeddings, row-parallel biases, MoE tensors) is replicated

The model geometry (heads, kv heads, SwiGLU) is read from the arguments stored in the checkpoint, or given
on the command line.
"""
from __future__ import annotations

import argparse
import re
import shutil
from pathlib import Path

import torch


class Geometry:
    def __init__(self, heads: int, kv_heads: int | None, swiglu: bool):
        self.heads, self.kv_heads, self.swiglu = heads, kv_heads or heads, swiglu


def _leaf(key: str) -> str:
    return key.split(".")[-1]


def _is_head_weight(key: str, layer_file: bool) -> bool:
    return key == "head.weight" or (layer_file and key == "weight")


def merge(key: str, shards: list[torch.Tensor], g: Geometry, layer_file: bool = False) -> torch.Tensor:
    """One full tensor out of the TP shards of `key`."""
    if len(shards) == 1 or all(torch.equal(shards[0], s) for s in shards[1:]):
        return shards[0].clone()
    leaf = _leaf(key)
    if leaf in ("wqkv", "bqkv"):
        n = len(shards)
        per = shards[0].shape[0] // (g.heads // n + 2 * (g.kv_heads // n))          # head_dim
        hq, hk = g.heads // n * per, g.kv_heads // n * per
        q = torch.cat([s[:hq] for s in shards])
        k = torch.cat([s[hq:hq + hk] for s in shards])
        v = torch.cat([s[hq + hk:] for s in shards])
        return torch.cat([q, k, v])
    if leaf in ("w1", "b1"):
        if g.swiglu:
            half = shards[0].shape[0] // 2
            return torch.cat([torch.cat([s[:half] for s in shards]), torch.cat([s[half:] for s in shards])])
        return torch.cat(shards)
    if leaf in ("wo", "w2"):
        return torch.cat(shards, dim=1)
    if _is_head_weight(key, layer_file):
        return torch.cat(shards)
    raise ValueError(f"{key}: shards differ but the tensor has no tensor-parallel rule")


def split(key: str, full: torch.Tensor, n: int, g: Geometry, layer_file: bool = False,
          head_replicated: bool = False) -> list[torch.Tensor]:
    """The `n` TP shards of a full tensor."""
    leaf = _leaf(key)
    if n == 1:
        return [full.clone()]
    if leaf in ("wqkv", "bqkv"):
        per = full.shape[0] // (g.heads + 2 * g.kv_heads)
        hq, hk = g.heads * per, g.kv_heads * per
        q, k, v = full[:hq], full[hq:hq + hk], full[hq + hk:]
        return [torch.cat([c.chunk(n)[r] for c in (q, k, v)]).clone() for r in range(n)]
    if leaf in ("w1", "b1") and "moe" not in key:
        if g.swiglu:
            half = full.shape[0] // 2
            return [torch.cat([full[:half].chunk(n)[r], full[half:].chunk(n)[r]]).clone() for r in range(n)]
        return [c.clone() for c in full.chunk(n)]
    if leaf in ("wo", "w2") and "moe" not in key:
        return [c.clone() for c in full.chunk(n, dim=1)]
    if _is_head_weight(key, layer_file) and not head_replicated:
        return [c.clone() for c in full.chunk(n)]
    return [full.clone() for _ in range(n)]


def _convert_state_dicts(sds: list[dict], target: int, g: Geometry, layer_file: bool) -> list[dict]:
    out = [dict() for _ in range(target)]
    for key in sds[0]:
        vals = [sd[key] for sd in sds]
        if not isinstance(vals[0], torch.Tensor):
            for o in out:
                o[key] = vals[0]
            continue
        replicated_head = _is_head_weight(key, layer_file) and len(vals) > 1 and all(torch.equal(vals[0], v) for v in vals[1:])
        full = merge(key, vals, g, layer_file)
        tied_single = _is_head_weight(key, layer_file) and len(vals) == 1 and not layer_file and \
            "embed.wte" in sds[0] and sds[0]["embed.wte"].shape == vals[0].shape and torch.equal(sds[0]["embed.wte"], vals[0])
        for r, piece in enumerate(split(key, full, target, g, layer_file, replicated_head or tied_single)):
            out[r][key] = piece
    return out


def reshard(load_dir: str, save_dir: str, target_tp: int, tag: str | None = None, geometry: Geometry | None = None) -> str:
    src_root, dst_root = Path(load_dir), Path(save_dir)
    tag = tag or (src_root / "latest").read_text().strip()
    src, dst = src_root / tag, dst_root / tag
    dst.mkdir(parents=True, exist_ok=True)
    mp_files = sorted(src.glob("mp_rank_*_model_states.pt"))
    assert mp_files, f"no mp_rank_*_model_states.pt under {src}"
    clients = [torch.load(f, map_location="cpu", weights_only=False) for f in mp_files]
    src_tp = len(clients)
    if geometry is None:
        a = clients[0].get("args") or {}
        assert a.get("num_attention_heads"), "model geometry not in the checkpoint: pass --heads / --kv-heads / --swiglu"
        geometry = Geometry(int(a["num_attention_heads"]), a.get("num_key_value_heads"), bool(a.get("swiglu")))
    g = geometry
    assert g.heads % target_tp == 0 and g.kv_heads % target_tp == 0, "target TP must divide the head counts"
    # whole-module files
    modules = [c.get("module") for c in clients]
    new_modules = _convert_state_dicts(modules, target_tp, g, False) if modules[0] is not None else [None] * target_tp
    for r in range(target_tp):
        c = {k: v for k, v in clients[0].items() if k != "module"}
        c["module"] = new_modules[r]
        par = dict(c.get("parallel") or {})
        par["tp"] = target_tp
        c["parallel"] = par
        if isinstance(c.get("args"), dict):
            c["args"] = dict(c["args"], tensor_model_parallel_size=target_tp)
        torch.save(c, dst / f"mp_rank_{r:02d}_model_states.pt")
    # per-layer files (pipeline layout)
    layers = sorted({int(m.group(1)) for f in src.glob("layer_*-model_*-model_states.pt")
                     if (m := re.match(r"layer_(\d+)-model_\d+-model_states\.pt", f.name))})
    for L in layers:
        sds = [torch.load(src / f"layer_{L:02d}-model_{t:02d}-model_states.pt", map_location="cpu", weights_only=False)
               for t in range(src_tp)]
        for r, sd in enumerate(_convert_state_dicts(sds, target_tp, g, True)):
            torch.save(sd, dst / f"layer_{L:02d}-model_{r:02d}-model_states.pt")
    # MoE expert files carry no TP sharding here (MoE layers run at TP = 1): copy them
    for f in src.glob("layer_*_expert_*_model_states.pt"):
        shutil.copy2(f, dst / f.name)
    (dst_root / "latest").write_text(tag)
    it = src_root / "latest_checkpointed_iteration.txt"
    if it.exists():
        shutil.copy2(it, dst_root / it.name)
    return str(dst)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="b200flow-reshard")
    ap.add_argument("--load", required=True)
    ap.add_argument("--save", required=True)
    ap.add_argument("--target-tp", type=int, required=True)
    ap.add_argument("--tag", default=None)
    ap.add_argument("--heads", type=int, default=None)
    ap.add_argument("--kv-heads", type=int, default=None)
    ap.add_argument("--swiglu", action="store_true")
    a = ap.parse_args(argv)
    geo = Geometry(a.heads, a.kv_heads, a.swiglu) if a.heads else None
    out = reshard(a.load, a.save, a.target_tp, a.tag, geo)
    print(f"wrote {out} with tensor-parallel size {a.target_tp}; load it with --no-load-optim")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
