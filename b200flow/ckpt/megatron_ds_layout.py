"""Megatron-DeepSpeed checkpoint directory layout.

The reference saves with `--save $CKPT_ROOT/$PET_NODE_RANK --save-interval 10000`
(examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-zero1.yaml:26,55,83).  File names kept
([UP] Megatron-DeepSpeed `save_checkpoint`):

    <save>/latest                                  text: global_step<N>
    <save>/latest_checkpointed_iteration.txt       text: <N>
    <save>/global_step<N>/
        mp_rank_<TP:02d>_model_states.pt                       module (non-pipeline) + client state
        layer_<idx:02d>-model_<TP:02d>-model_states.pt         per layer (pipeline parallel)
        zero_pp_rank_<DP>_mp_rank_<TP:02d>_optim_states.pt     ZeRO-1 shard (fp16)
        bf16_zero_pp_rank_<DP>_mp_rank_<TP:02d>_optim_states.pt   … under bf16
        layer_<L>_expert_<E>_mp_rank_<TP:02d>_model_states.pt  MoE expert weights
        expp_rank_<EP>_mp_rank_<TP:02d>_optim_states.pt        MoE expert optimizer shard (expert-DP rank 0)
        expp_rank_<EP>_edp_<R>_mp_rank_<TP:02d>_optim_states.pt   … of expert-data-parallel rank R > 0

Pipeline layer indices follow the PipelineModule convention: embedding = 00, transformer
layer i = i + 1, final norm + head = layers + 1.  Saves run on a background thread
(`async_save=True`) once tensors are staged to host, since a 7B model with optimizer state
is ~100 GB.
"""
from __future__ import annotations

import threading
from pathlib import Path

import torch
import torch.distributed as dist

_pending: list[threading.Thread] = []


def _cpu(sd):
    return {k: (v.detach().to("cpu", copy=True) if isinstance(v, torch.Tensor) else v) for k, v in sd.items()}


def _write(obj, path: Path, async_save: bool):
    if async_save:
        t = threading.Thread(target=torch.save, args=(obj, path), daemon=False)
        t.start()
        _pending.append(t)
    else:
        torch.save(obj, path)


def wait_for_pending_saves() -> None:
    while _pending:
        _pending.pop().join()


def tag_for(step: int) -> str:
    return f"global_step{step}"


def layer_file(idx: int, tp: int) -> str:
    return f"layer_{idx:02d}-model_{tp:02d}-model_states.pt"


def model_states_file(tp: int) -> str:
    return f"mp_rank_{tp:02d}_model_states.pt"


def optim_file(dp: int, tp: int, bf16: bool) -> str:
    return ("bf16_" if bf16 else "") + f"zero_pp_rank_{dp}_mp_rank_{tp:02d}_optim_states.pt"


def expert_file(layer: int, expert: int, tp: int) -> str:
    return f"layer_{layer}_expert_{expert}_mp_rank_{tp:02d}_model_states.pt"


def expert_optim_file(ep: int, tp: int, edp: int = 0) -> str:
    """The expert optimizer is ZeRO-partitioned over the expert-data-parallel group (dp / ep ranks that hold the
    same experts), so every (EP rank, expert-DP rank) pair owns a distinct shard and needs its own file;
    expert-DP rank 0 keeps DeepSpeed's name."""
    mid = f"_edp_{edp}" if edp else ""
    return f"expp_rank_{ep}{mid}_mp_rank_{tp:02d}_optim_states.pt"


def _edp_rank(engine) -> int:
    eo = getattr(engine, "expert_optimizer", None)
    return int(getattr(eo, "dp_rank", 0)) if eo is not None else 0


_EXPERT_TENSORS = ("w1", "b1", "w2", "b2")     # MoELayer keeps its local experts batched: [E_local, ...]


def _split_expert_state(sd: dict, ep_rank: int = 0):
    """dense state dict + {(layer, global expert id): {key: that expert's slice}}."""
    dense, experts = {}, {}
    for k, v in sd.items():
        parts = k.split(".")
        if len(parts) >= 2 and parts[-2] == "moe" and parts[-1] in _EXPERT_TENSORS and v.dim() >= 2:
            L = int(parts[parts.index("layers") + 1]) if "layers" in parts else 0
            for e in range(v.shape[0]):
                experts.setdefault((L, ep_rank * v.shape[0] + e), {})[k] = v[e].clone()
        else:
            dense[k] = v
    return dense, experts


def _ep_rank(engine) -> int:
    ctx = getattr(engine, "ep_ctx", None)
    return int(getattr(ctx, "rank", 0)) if ctx is not None else 0


def save_checkpoint(engine, save_dir: str, tag: str | None = None, client_state: dict | None = None,
                    async_save: bool = False, save_optimizer: bool = True, save_rng: bool = True) -> str:
    ps = engine.ps
    tp = ps.tp_rank if ps else 0
    dp = ps.dp_rank if ps else 0
    pp_world = engine.pp_world
    step = engine.global_steps
    tag = tag or tag_for(step)
    d = Path(save_dir) / tag
    d.mkdir(parents=True, exist_ok=True)
    bf16 = engine.dtype == torch.bfloat16
    model = engine.model
    rng = {}
    if save_rng:     # CPU + CUDA generators and the engine's dropout-call counter: masks continue, not restart
        rng = {"rng_state": torch.get_rng_state(), "drop_calls": int(getattr(engine, "_drop_calls", 0))}
        if torch.cuda.is_available() and engine.device.type == "cuda":
            rng["cuda_rng_state"] = torch.cuda.get_rng_state(engine.device)
    client = {"iteration": step, "global_steps": step, "args": getattr(engine, "args", None),
              **rng, "lr": engine.optimizer.lr,
              "parallel": {"tp": engine.tp_world, "pp": pp_world, "dp": engine.dp_world},
              **(client_state or {})}
    if dp == 0:      # one DP replica writes the module
        if pp_world > 1:
            lo, hi = model.layer_range
            if model.embed is not None:
                _write(_cpu(model.embed.state_dict()), d / layer_file(0, tp), async_save)
            for i, layer in zip(range(lo, hi), model.layers):
                _write(_cpu(layer.state_dict()), d / layer_file(i + 1, tp), async_save)
            if model.head is not None:
                _write(_cpu(model.head.state_dict()), d / layer_file(engine.cfg.layers + 1, tp), async_save)
            if engine.stage == 0:
                _write({"module": None, **client}, d / model_states_file(tp), async_save)
        else:
            dense, experts = _split_expert_state(_cpu(model.state_dict()), _ep_rank(engine))
            _write({"module": dense, **client}, d / model_states_file(tp), async_save)
            for (L, E), sd in experts.items():     # one file per (layer, global expert), as DeepSpeed MoE does
                _write(sd, d / expert_file(L, E, tp), async_save)
    elif getattr(engine, "ep_world", 1) > 1 and dp < engine.ep_world and pp_world == 1:
        # the other EP ranks of the first expert-data-parallel replica own the remaining experts
        _, experts = _split_expert_state(_cpu(model.state_dict()), _ep_rank(engine))
        for (L, E), sd in experts.items():
            _write(sd, d / expert_file(L, E, tp), async_save)
    if save_optimizer:        # Megatron's --no-save-optim leaves the ZeRO partitions out (load with --no-load-optim)
        osd = engine.optimizer.state_dict()
        osd = {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in osd.items()}
        osd["stage"] = engine.stage
        scaler = getattr(engine, "loss_scaler", None)
        if scaler is not None:     # DeepSpeed keeps the fp16 loss-scale state with the optimizer state
            osd["loss_scaler"] = scaler.state_dict()
            osd["skipped_steps"] = engine.skipped_steps
        name = optim_file(dp, tp, bf16)
        if pp_world > 1:   # one file per (stage, dp, tp): stage goes into the DeepSpeed "pp_rank" slot
            name = ("bf16_" if bf16 else "") + f"zero_pp_rank_{dp + engine.dp_world * engine.stage}_mp_rank_{tp:02d}_optim_states.pt"
        _write({"optimizer_state_dict": osd, "ds_version": "b200flow"}, d / name, async_save)
        eopt = getattr(engine, "expert_optimizer", None)
        if eopt is not None:        # expert parameters have their own (expert-data-parallel) optimizer shard
            esd = {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in eopt.state_dict().items()}
            _write({"optimizer_state_dict": esd, "ds_version": "b200flow"},
                   d / expert_optim_file(_ep_rank(engine), tp, _edp_rank(engine)), async_save)
    # `latest` may only name a complete checkpoint.  Every rank names the checkpoint in ITS save directory: the
    # reference's examples save to `$CKPT_ROOT/$PET_NODE_RANK`, one directory per node, and a node without
    # `latest` could not resume.  Same content from every writer, put in place by a rename.
    import os

    def publish():
        for name, text in (("latest", tag), ("latest_checkpointed_iteration.txt", str(step))):
            tmp = Path(save_dir) / f".{name}.{os.getpid()}.tmp"
            tmp.write_text(text)
            os.replace(tmp, Path(save_dir) / name)

    if not async_save:
        wait_for_pending_saves()             # earlier asynchronous saves, if any
        if dist.is_initialized():
            dist.barrier()                   # every rank's files are on disk
        publish()
        return str(d)
    # asynchronous: training continues now; a publisher thread joins this rank's writers, drops a marker into the
    # checkpoint directory and renames `latest` into place once the markers of every rank that writes into this
    # directory are there - a crash before that leaves `latest` on the previous, complete checkpoint
    rank = dist.get_rank() if dist.is_initialized() else 0
    peers = 1
    if dist.is_initialized():
        dirs = [None] * dist.get_world_size()
        dist.all_gather_object(dirs, str(Path(save_dir).resolve()))
        peers = sum(1 for x in dirs if x == str(Path(save_dir).resolve()))
    writers = list(_pending)
    del _pending[:]

    def finish():
        import time

        for t in writers:
            t.join()
        (d / f".complete_rank_{rank}").write_text(tag)
        deadline = time.monotonic() + 3600
        while len(list(d.glob(".complete_rank_*"))) < peers and time.monotonic() < deadline:
            time.sleep(0.05)
        if len(list(d.glob(".complete_rank_*"))) >= peers:
            publish()

    t = threading.Thread(target=finish, daemon=False)
    t.start()
    _pending.append(t)
    return str(d)


def _optim_name(engine, dp: int, dp_world: int, tp: int, bf16: bool) -> str:
    if engine.pp_world > 1:
        return ("bf16_" if bf16 else "") + f"zero_pp_rank_{dp + dp_world * engine.stage}_mp_rank_{tp:02d}_optim_states.pt"
    return optim_file(dp, tp, bf16)


def _load_optimizer_state(d: Path, name: str, engine, tp: int, bf16: bool) -> dict:
    """This rank's ZeRO-1 optimizer state.  When the checkpoint was written by a different data-parallel size
    (a job restarted on fewer or more GPUs — DeepSpeed's elastic checkpoint for ZeRO-1), the saved partitions of
    this (stage, TP rank) are joined into the full flat state and cut again for the current layout; only the
    zero padding at the end of the flat buffer can differ in length."""
    opt = engine.optimizer
    own = d / name
    if own.exists():
        osd = torch.load(own, map_location="cpu", weights_only=False)["optimizer_state_dict"]
        if int(osd.get("dp_world", engine.dp_world)) == engine.dp_world and tuple(osd["partition"]) == (opt.lo, opt.hi):
            return osd
    first = d / _optim_name(engine, 0, 1, tp, bf16)      # dp index 0 has the same name for every saved dp size
    if engine.pp_world > 1:
        # under PP the file index mixes stage and dp: find the saved dp size from any file of this stage
        first = None
        for cand in range(1, 65):
            f = d / _optim_name(engine, 0, cand, tp, bf16)
            if f.exists():
                o = torch.load(f, map_location="cpu", weights_only=False)["optimizer_state_dict"]
                if int(o.get("stage", engine.stage)) == engine.stage and int(o.get("dp_world", cand)) == cand:
                    first = f
                    break
        assert first is not None, f"no optimizer state for stage {engine.stage}, TP rank {tp} in {d}"
    head = torch.load(first, map_location="cpu", weights_only=False)["optimizer_state_dict"]
    saved_dp = int(head["dp_world"])
    parts = [head] + [torch.load(d / _optim_name(engine, i, saved_dp, tp, bf16), map_location="cpu",
                                 weights_only=False)["optimizer_state_dict"] for i in range(1, saved_dp)]
    parts.sort(key=lambda o: o["partition"][0])
    out = dict(head)
    for key in ("fp32_flat", "exp_avg", "exp_avg_sq"):
        full = torch.cat([o[key].reshape(-1) for o in parts])
        now = torch.zeros(engine.flat.numel, dtype=full.dtype)
        n = min(full.numel(), now.numel())
        now[:n] = full[:n]
        out[key] = now[opt.lo:opt.hi].clone()
    out["partition"], out["dp_world"], out["dp_rank"] = (opt.lo, opt.hi), engine.dp_world, opt.dp_rank
    return out


def _restore_rng(engine, client: dict) -> None:
    if "rng_state" in client:
        torch.set_rng_state(client["rng_state"])
    if "cuda_rng_state" in client and engine.device.type == "cuda":
        torch.cuda.set_rng_state(client["cuda_rng_state"], engine.device)
    if "drop_calls" in client and hasattr(engine, "_drop_calls"):
        engine._drop_calls = int(client["drop_calls"])


def load_checkpoint(engine, load_dir: str, tag: str | None = None, load_optimizer: bool = True,
                    load_rng: bool = True, finetune: bool = False):
    """`load_optimizer=False` / `load_rng=False` / `finetune=True` are Megatron's `--no-load-optim`,
    `--no-load-rng` and `--finetune` (weights only, iteration back to 0)."""
    wait_for_pending_saves()
    root = Path(load_dir)
    if tag is None:
        latest = root / "latest"
        if not latest.exists():
            return None, {}
        tag = latest.read_text().strip()
    d = root / tag
    ps = engine.ps
    tp = ps.tp_rank if ps else 0
    dp = ps.dp_rank if ps else 0
    model = engine.model
    bf16 = engine.dtype == torch.bfloat16
    client = {}
    if engine.pp_world > 1:
        lo, hi = model.layer_range
        if model.embed is not None:
            model.embed.load_state_dict(torch.load(d / layer_file(0, tp), map_location="cpu"))
        for i, layer in zip(range(lo, hi), model.layers):
            layer.load_state_dict(torch.load(d / layer_file(i + 1, tp), map_location="cpu"))
        if model.head is not None:
            model.head.load_state_dict(torch.load(d / layer_file(engine.cfg.layers + 1, tp), map_location="cpu"),
                                       strict=False)
        ms = d / model_states_file(tp)
        if ms.exists():
            client = torch.load(ms, map_location="cpu", weights_only=False)
        name = ("bf16_" if bf16 else "") + f"zero_pp_rank_{dp + engine.dp_world * engine.stage}_mp_rank_{tp:02d}_optim_states.pt"
    else:
        client = torch.load(d / model_states_file(tp), map_location="cpu", weights_only=False)
        sd = dict(client.pop("module"))
        own = model.state_dict()
        ep_rank = _ep_rank(engine)
        for k, v in own.items():     # restack this rank's experts from the per-expert files
            parts = k.split(".")
            if len(parts) >= 2 and parts[-2] == "moe" and parts[-1] in _EXPERT_TENSORS and v.dim() >= 2:
                L = int(parts[parts.index("layers") + 1]) if "layers" in parts else 0
                rows = [torch.load(d / expert_file(L, ep_rank * v.shape[0] + e, tp), map_location="cpu")[k]
                        for e in range(v.shape[0])]
                sd[k] = torch.stack(rows)
        model.load_state_dict(sd, strict=False)
        name = optim_file(dp, tp, bf16)
    if not load_optimizer or finetune:      # weights only: master copies follow the loaded weights, moments stay zero
        for o in (engine.optimizer, getattr(engine, "expert_optimizer", None)):
            if o is not None:
                o.master.copy_(o.flat.params[o.lo:o.hi].float())
        engine.global_steps = 0 if finetune else int(client.get("global_steps", 0))
        if load_rng and not finetune:
            _restore_rng(engine, client)
        return str(d), client
    osd = _load_optimizer_state(d, name, engine, tp, bf16)
    dev = engine.flat.params.device
    osd = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in osd.items()}
    engine.optimizer.load_state_dict(osd)
    if getattr(engine, "loss_scaler", None) is not None and "loss_scaler" in osd:
        engine.loss_scaler.load_state_dict(osd["loss_scaler"])
        engine.skipped_steps = int(osd.get("skipped_steps", 0))
    eopt = getattr(engine, "expert_optimizer", None)
    ef = d / expert_optim_file(_ep_rank(engine), tp, _edp_rank(engine))
    if eopt is not None and ef.exists():
        esd = torch.load(ef, map_location="cpu", weights_only=False)["optimizer_state_dict"]
        eopt.load_state_dict({k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in esd.items()})
    engine.global_steps = int(client.get("global_steps", osd.get("step", 0)))
    if load_rng:
        _restore_rng(engine, client)
    return str(d), client
