"""Megatron-DeepSpeed checkpoint layout round trips on CPU: file names, `latest`, and a
restored engine that continues exactly like the one that saved — dense GPT and GPT-MoE (expert files)."""
import itertools

import pytest
import torch

from b200flow.engines.ds_engine import initialize
from b200flow.models.gpt import config_from_preset


def _engine(preset, seed=11):
    cfg = config_from_preset(preset)
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1,
                           "optimizer": {"params": {"lr": 1e-3}}}, device=torch.device("cpu"), seed=seed)
    g = torch.Generator().manual_seed(5)
    toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(3)]
    batches = [(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks]
    return cfg, eng, batches


@pytest.mark.parametrize("preset,expert_files", [("gpt2-tiny", False), ("gpt-moe-tiny", True)])
def test_save_load_continues_identically(tmp_path, preset, expert_files):
    cfg, eng, batches = _engine(preset)
    it = itertools.cycle(batches)
    for _ in range(2):
        eng.train_batch(it)
    eng.save_checkpoint(str(tmp_path))
    assert (tmp_path / "latest").read_text() == "global_step2"
    assert (tmp_path / "latest_checkpointed_iteration.txt").read_text().strip() == "2"
    names = {p.name for p in (tmp_path / "global_step2").iterdir()}
    assert "mp_rank_00_model_states.pt" in names
    assert any(n.endswith("zero_pp_rank_0_mp_rank_00_optim_states.pt") for n in names)
    if expert_files:
        assert any(n.startswith("layer_") and "_expert_0_mp_rank_00_model_states.pt" in n for n in names), names
        assert "expp_rank_0_mp_rank_00_optim_states.pt" in names
    after_save = [float(eng.train_batch(it)) for _ in range(2)]

    _, fresh, _ = _engine(preset, seed=99)          # different init: everything must come from the files
    path, client = fresh.load_checkpoint(str(tmp_path))
    assert fresh.global_steps == 2 and path.endswith("global_step2")
    it2 = itertools.cycle(batches)
    for _ in range(2):
        next(it2)                                    # the data position is the caller's business
    resumed = [float(fresh.train_batch(it2)) for _ in range(2)]
    assert resumed == pytest.approx(after_save, rel=1e-5, abs=1e-6)


def _dp2_worker(rank, port, save_dir, mode, q, world=2, pp=1):
    import os

    import torch.distributed as dist

    from b200flow.parallel.groups import ParallelState

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    ps = ParallelState.initialize(tp=1, pp=pp)
    cfg = config_from_preset("gpt2-tiny")
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1,
                           "optimizer": {"params": {"lr": 1e-3}}}, ps, device=torch.device("cpu"), seed=11)
    if mode == "save":
        g = torch.Generator().manual_seed(5 + ps.dp_rank)
        toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(3)]
        it = itertools.cycle([(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks])
        for _ in range(3):
            eng.train_batch(it)
        eng.save_checkpoint(save_dir)
        o = eng.optimizer
        q.put((rank, o.lo, o.hi, eng.flat.params.numpy().copy(), o.exp_avg.numpy().copy(), o.exp_avg_sq.numpy().copy(),
               o.step_count, eng.stage, ps.dp_rank))       # numpy: pickled by value, safe after this process exits
    else:
        eng.load_checkpoint(save_dir)
        o = eng.optimizer
        q.put((rank, o.lo, o.hi, eng.flat.params.numpy().copy(), o.master.numpy().copy(), o.exp_avg.numpy().copy(),
               o.step_count, eng.stage, ps.dp_rank))
    dist.barrier()
    dist.destroy_process_group()


def _spawn(mode, save_dir, port, world=2, pp=1):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_dp2_worker, args=(r, port, save_dir, mode, q, world, pp)) for r in range(world)]
    for p in procs:
        p.start()
    import numpy as np

    out = sorted((tuple(torch.from_numpy(v) if isinstance(v, np.ndarray) else v for v in q.get(timeout=300))
                  for _ in procs), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return out


def test_zero1_state_is_resharded_when_the_data_parallel_size_changes(tmp_path):
    """A job saved on 2 data-parallel ranks resumes on 1, and the other way round (elastic restart on a
    different number of GPUs): the partitions are joined and cut again."""
    two = _spawn("save", str(tmp_path / "dp2"), 29681)
    names = {p.name for p in (tmp_path / "dp2" / "global_step3").iterdir()}
    assert {n for n in names if "optim_states" in n} == {"zero_pp_rank_0_mp_rank_00_optim_states.pt",
                                                         "zero_pp_rank_1_mp_rank_00_optim_states.pt"}
    _, one, _ = _engine("gpt2-tiny", seed=99)
    one.load_checkpoint(str(tmp_path / "dp2"))
    o = one.optimizer
    n = min(one.flat.numel, two[0][3].numel())
    assert torch.equal(one.flat.params[:n], two[0][3][:n]) and torch.equal(o.master[:n], two[0][3][:n].float())
    joined = torch.cat([two[0][4], two[1][4]])
    assert torch.equal(o.exp_avg[:n], joined[:n]) and o.exp_avg.abs().sum() > 0 and o.step_count == 3
    assert one.global_steps == 3
    # 1 -> 2: save from the single-process engine, load on two ranks
    one.save_checkpoint(str(tmp_path / "dp1"))
    back = _spawn("load", str(tmp_path / "dp1"), 29682)
    for rank, lo, hi, params, master, exp_avg, steps, _stage, _dp in back:
        m = min(n, hi)
        assert torch.equal(master[: m - lo], one.flat.params[lo:m].float())
        assert torch.equal(exp_avg[: m - lo], o.exp_avg[lo:m]) and steps == 3
        assert torch.equal(params[:n], one.flat.params[:n])


def test_zero1_resharding_under_pipeline_parallelism(tmp_path):
    """PP2 x DP2 (4 ranks) -> PP2 x DP1 (2 ranks): each stage joins the two partitions of its own layers."""
    saved = _spawn("save", str(tmp_path / "pp2dp2"), 29683, world=4, pp=2)
    loaded = _spawn("load", str(tmp_path / "pp2dp2"), 29684, world=2, pp=2)
    for rank, lo, hi, params, master, exp_avg, steps, stage, _dp in loaded:
        mine = sorted((t for t in saved if t[7] == stage), key=lambda t: t[1])        # the stage's two partitions
        assert len(mine) == 2 and steps == 3
        joined = torch.cat([mine[0][4], mine[1][4]])
        n = min(joined.numel(), exp_avg.numel())
        assert torch.equal(exp_avg[:n], joined[:n]) and exp_avg.abs().sum() > 0
        assert torch.equal(params[:n], mine[0][3][:n]) and torch.equal(master[:n], params[:n].float())


def test_finetune_and_no_load_optim_take_weights_only(tmp_path):
    cfg, eng, batches = _engine("gpt2-tiny")
    it = itertools.cycle(batches)
    for _ in range(2):
        eng.train_batch(it)
    eng.save_checkpoint(str(tmp_path))
    _, a, _ = _engine("gpt2-tiny", seed=99)
    a.load_checkpoint(str(tmp_path), load_optimizer_states=False)
    assert torch.equal(a.flat.params, eng.flat.params) and a.global_steps == 2
    assert float(a.optimizer.exp_avg.abs().sum()) == 0.0 and a.optimizer.step_count == 0
    assert torch.equal(a.optimizer.master, a.flat.params[a.optimizer.lo:a.optimizer.hi].float())
    _, b, _ = _engine("gpt2-tiny", seed=98)
    b.load_checkpoint(str(tmp_path), finetune=True)
    assert torch.equal(b.flat.params, eng.flat.params) and b.global_steps == 0
    assert float(b.train_batch(itertools.cycle(batches))) > 0


def test_no_save_optim_writes_weights_only(tmp_path):
    cfg, eng, batches = _engine("gpt2-tiny")
    eng.train_batch(itertools.cycle(batches))
    eng.save_checkpoint(str(tmp_path), save_optimizer=False, save_rng=False)
    names = {p.name for p in (tmp_path / "global_step1").iterdir()}
    assert "mp_rank_00_model_states.pt" in names and not any("optim_states" in n for n in names)
    client = torch.load(tmp_path / "global_step1" / "mp_rank_00_model_states.pt", weights_only=False)
    assert "rng_state" not in client
    _, fresh, _ = _engine("gpt2-tiny", seed=99)
    fresh.load_checkpoint(str(tmp_path), load_optimizer_states=False)
    assert torch.equal(fresh.flat.params, eng.flat.params)


def _moe_ckpt_worker(rank, world, ep, port, save_dir, q):
    """GPT-MoE engine at data-parallel size `world`, expert-parallel size `ep`: save, keep training, then a fresh
    engine loads and must continue identically (every (EP rank, expert-DP rank) shard in its own file)."""
    import os

    import torch.distributed as dist

    from b200flow.parallel.groups import ParallelState

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    ps = ParallelState.initialize(tp=1, pp=1, ep=ep)
    cfg = config_from_preset("gpt-moe-tiny")
    dsc = {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1, "optimizer": {"params": {"lr": 1e-3}}}
    g = torch.Generator().manual_seed(5 + rank)
    toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(4)]
    batches = [(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks]
    eng = initialize(cfg, dsc, ps, device=torch.device("cpu"), seed=11)
    it = itertools.cycle(batches)
    for _ in range(2):
        eng.train_batch(it)
    eng.save_checkpoint(save_dir)
    after = [float(eng.train_batch(it)) for _ in range(2)]
    fresh = initialize(cfg, dsc, ps, device=torch.device("cpu"), seed=99)
    fresh.load_checkpoint(save_dir)
    it2 = itertools.cycle(batches)
    next(it2), next(it2)
    resumed = [float(fresh.train_batch(it2)) for _ in range(2)]
    eo = eng.expert_optimizer
    q.put((rank, after, resumed, (eo.lo, eo.hi), eo.dp_rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,ep,port", [(4, 2, 29687), (2, 1, 29688)])
def test_expert_optimizer_shards_do_not_collide_when_dp_exceeds_ep(tmp_path, world, ep, port):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_moe_ckpt_worker, args=(r, world, ep, port, str(tmp_path), q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    names = {p.name for p in (tmp_path / "global_step2").iterdir() if p.name.startswith("expp_rank_")}
    assert len(names) == world, names            # one expert optimizer shard per rank: (ep ranks) x (expert-dp ranks)
    assert "expp_rank_0_mp_rank_00_optim_states.pt" in names
    for rank, after, resumed, _part, _edp in out:
        assert resumed == pytest.approx(after, rel=1e-5, abs=1e-6), (rank, after, resumed)


def test_async_save_publishes_latest_only_when_the_files_are_complete(tmp_path, monkeypatch):
    from b200flow.ckpt import megatron_ds_layout as L

    cfg, eng, batches = _engine("gpt2-tiny")
    it = itertools.cycle(batches)
    eng.train_batch(it)
    eng.save_checkpoint(str(tmp_path))                           # step 1, synchronous
    eng.train_batch(it)
    import threading

    gate = threading.Event()
    real_save = torch.save

    def slow_save(obj, path, *a, **kw):
        gate.wait(timeout=60)
        return real_save(obj, path, *a, **kw)

    monkeypatch.setattr(torch, "save", slow_save)
    eng.save_checkpoint(str(tmp_path), async_save=True)         # returns while the writers are still blocked
    assert (tmp_path / "latest").read_text() == "global_step1"   # a crash here resumes from the complete checkpoint
    gate.set()
    L.wait_for_pending_saves()
    monkeypatch.setattr(torch, "save", real_save)
    assert (tmp_path / "latest").read_text() == "global_step2"
    _, fresh, _ = _engine("gpt2-tiny", seed=99)
    fresh.load_checkpoint(str(tmp_path))
    assert fresh.global_steps == 2 and torch.equal(fresh.flat.params, eng.flat.params)


def test_resume_with_dropout_reproduces_the_masks(tmp_path):
    """The dropout-call counter travels with the checkpoint, so the resumed run draws the masks the
    uninterrupted run draws (not the ones from step 0 again)."""
    def make(seed):
        cfg = config_from_preset("gpt2-tiny")
        cfg.dropout = 0.1
        return cfg, initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1,
                                     "optimizer": {"params": {"lr": 1e-3}}}, device=torch.device("cpu"), seed=seed)

    cfg, eng = make(11)
    g = torch.Generator().manual_seed(5)
    toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(3)]
    batches = [(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks]
    it = itertools.cycle(batches)
    for _ in range(2):
        eng.train_batch(it)
    eng.save_checkpoint(str(tmp_path))
    after = [float(eng.train_batch(it)) for _ in range(2)]
    _, fresh = make(11)
    fresh.load_checkpoint(str(tmp_path))
    assert fresh._drop_calls == 2
    it2 = itertools.cycle(batches)
    next(it2), next(it2)
    resumed = [float(fresh.train_batch(it2)) for _ in range(2)]
    assert resumed == pytest.approx(after, rel=1e-5, abs=1e-6)
