"""Tensor-parallel resharding of Megatron-DeepSpeed-layout checkpoints (b200flow/ckpt/reshard.py): a model saved
at TP 2 evaluates identically after conversion to TP 1 and back, with and without pipeline parallelism."""
import itertools
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.ckpt.reshard import Geometry, merge, reshard, split  # noqa: E402

GEO = Geometry(heads=4, kv_heads=2, swiglu=True)


def _engine(ps):
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, kv_heads=2, ffn=128, layers=4, vocab_size=128, max_seq=32)
    return cfg, initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
                                 "optimizer": {"params": {"lr": 2e-3}}}, ps, device=torch.device("cpu"), seed=21)


def _batches(cfg):
    g = torch.Generator().manual_seed(0)
    mbs = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(4)]
    return [(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in mbs]


def _worker(rank, world, tp, pp, port, mode, ckpt, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    ps = None
    if world > 1:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from b200flow.parallel.groups import ParallelState

        ps = ParallelState.initialize(tp=tp, pp=pp)
    torch.set_num_threads(1)
    cfg, eng = _engine(ps)
    data = _batches(cfg)
    if mode == "train_save":
        it = itertools.cycle(data)
        for _ in range(2):
            eng.train_batch(it)
        eng.save_checkpoint(ckpt)
    else:
        eng.load_checkpoint(ckpt, load_optimizer_states=False)
    ev = float(eng.eval_batch(itertools.cycle(data), num_micro_batches=2))
    if rank == 0:
        q.put(ev)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _run(world, tp, pp, port, mode, ckpt):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, tp, pp, port, mode, ckpt, q)) for r in range(world)]
    for p in procs:
        p.start()
    ev = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return ev


def test_merge_and_split_are_inverse_for_every_rule():
    torch.manual_seed(0)
    d = 16
    full = {"layers.0.wqkv": torch.randn((4 + 2 * 2) * d, 64), "layers.0.bqkv": torch.randn((4 + 2 * 2) * d),
            "layers.0.w1": torch.randn(2 * 128, 64), "layers.0.wo": torch.randn(64, 4 * d),
            "layers.0.w2": torch.randn(64, 128), "head.weight": torch.randn(128, 64), "layers.0.ln1_w": torch.randn(64)}
    for key, t in full.items():
        shards = split(key, t, 2, GEO)
        assert torch.equal(merge(key, shards, GEO), t), key
    q_rows = split("layers.0.wqkv", full["layers.0.wqkv"], 2, GEO)[1][:2 * d]          # rank 1 owns query heads 2, 3
    assert torch.equal(q_rows, full["layers.0.wqkv"][2 * d:4 * d])
    assert all(torch.equal(s, full["layers.0.ln1_w"]) for s in split("layers.0.ln1_w", full["layers.0.ln1_w"], 2, GEO))


def test_tp2_checkpoint_converted_to_tp1_and_back_evaluates_the_same(tmp_path):
    a, b, c = (str(tmp_path / n) for n in ("tp2", "tp1", "tp2again"))
    want = _run(2, 2, 1, 29691, "train_save", a)
    reshard(a, b, 1, geometry=GEO)
    assert abs(_run(1, 1, 1, 29692, "load", b) - want) < 2e-4
    reshard(b, c, 2, geometry=GEO)
    assert abs(_run(2, 2, 1, 29693, "load", c) - want) < 2e-4


def test_resharding_the_per_layer_files_of_a_pipeline_checkpoint(tmp_path):
    a, b = str(tmp_path / "tp2pp2"), str(tmp_path / "tp1pp2")
    want = _run(4, 2, 2, 29694, "train_save", a)
    reshard(a, b, 1, geometry=GEO)
    names = {p.name for p in (Path(b) / "global_step2").iterdir()}
    assert "layer_01-model_00-model_states.pt" in names and not any("model_01" in n for n in names)
    assert abs(_run(2, 1, 2, 29695, "load", b) - want) < 2e-4


def _tied_worker(rank, world, port, q, pp=1):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    ps = None
    if world > 1:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from b200flow.parallel.groups import ParallelState

        ps = ParallelState.initialize(tp=world // pp, pp=pp)
    torch.set_num_threads(1)
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import GPTModel, config_from_preset

    tp_w, tp_r = (ps.topo.tp, ps.tp_rank) if ps else (1, 0)

    cfg = config_from_preset("gpt2-tiny", hidden=64, heads=4, ffn=128, vocab_size=128, max_seq=32)   # tied, biases, GELU
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
                           "optimizer": {"params": {"lr": 2e-3, "weight_decay": 0.0}}}, ps, device=torch.device("cpu"), seed=4)
    torch.manual_seed(77)
    full = GPTModel(cfg, dtype=torch.float32, device=torch.device("cpu")).state_dict()
    geo = Geometry(heads=4, kv_heads=4, swiglu=False)
    lo = eng.model.layer_range[0]
    with torch.no_grad():
        for name, p in eng.model.named_parameters():
            src = name
            if name.startswith("layers."):
                parts = name.split(".")
                src = ".".join(["layers", str(int(parts[1]) + lo)] + parts[2:])
            if name == "head.weight" and pp > 1:      # last stage: its own vocab-parallel copy of the tied table
                p.copy_(full["embed.wte"].chunk(tp_w)[tp_r])
                continue
            p.copy_(split(src, full[src], tp_w, geo, head_replicated=True)[tp_r])
    o = eng.optimizer
    o.master.copy_(eng.flat.params[o.lo:o.hi].float())
    it = itertools.cycle(_batches(cfg))
    losses = [float(eng.train_batch(it)) for _ in range(4)]
    same = True
    if world > 1 and pp == 1:
        wte = eng.model.embed.wte.detach().clone()
        other = [torch.empty_like(wte) for _ in range(world)]
        dist.all_gather(other, wte)
        same = all(torch.allclose(other[0], t, atol=1e-6) for t in other)
    if pp > 1:            # the last stage's head rows must still equal the first stage's table rows
        mine = eng.model.embed.wte.detach()[tp_r * (cfg.vocab_size // tp_w):(tp_r + 1) * (cfg.vocab_size // tp_w)] \
            if eng.stage == 0 else eng.model.head.weight.detach()
        pair = [torch.empty_like(mine) for _ in range(2)]
        dist.all_gather(pair, mine.contiguous(), group=ps.embedding_group)
        ok = torch.tensor([float(torch.allclose(pair[0], pair[1], atol=1e-6))])
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        same = bool(ok.item())
    if rank == 0:
        q.put((losses, same))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def test_tied_embeddings_under_tensor_parallelism_without_pipeline_stages():
    """GPT-2 (tied input / output embeddings) at TP 2, PP 1: the head works on each rank's vocabulary rows of the
    shared table, the table stays identical on both ranks, and the loss curve is the single-process one."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    _tied_worker(0, 1, 29696, q)
    want, _ = q.get()
    procs = [ctx.Process(target=_tied_worker, args=(r, 2, 29697, q)) for r in range(2)]
    for p in procs:
        p.start()
    got, same = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert same, "the replicated embedding table diverged between TP ranks"
    assert got[-1] < got[0]
    for a, b in zip(got, want):
        assert abs(a - b) < 2e-3, (got, want)


def test_tied_embeddings_under_tp2_pp2_follow_the_single_process_curve():
    """The reference's GPT-2 345M TP2 x PP2 example ties the output layer to the input embedding across the first
    and the last pipeline stage: same loss curve as one process, head rows stay equal to the
    table rows."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    _tied_worker(0, 1, 29698, q)
    want, _ = q.get()
    procs = [ctx.Process(target=_tied_worker, args=(r, 4, 29699, q, 2)) for r in range(4)]
    for p in procs:
        p.start()
    got, same = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert same, "tied head rows and embedding rows diverged"
    for a, b in zip(got, want):
        assert abs(a - b) < 2e-3, (got, want)
