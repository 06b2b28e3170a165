"""fp16 loss scaling (DeepSpeed's dynamic scaler semantics, skipped steps, checkpointed state) and
forward-only evaluation under pipeline parallelism, on CPU/gloo.

The reference's Megatron-DeepSpeed values files train with `--fp16` + `"fp16": {"enabled": true}` and
evaluate every `--eval-interval` on a TP2 x PP2 topology
(examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml:33-57)."""
import itertools
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from b200flow.parallel.loss_scaler import (DynamicLossScaler, LossScaler, LossScalerConfig,  # noqa: E402
                                           make_loss_scaler)


def test_dynamic_scaler_follows_deepspeed_state_machine():
    s = DynamicLossScaler(LossScalerConfig(initial_scale_power=4, scale_window=3, hysteresis=2, min_scale=1))
    assert s.scale == 16
    s.update(True)                      # first overflow only uses up hysteresis
    assert s.scale == 16 and s.cur_hysteresis == 1 and s.skipped == 1
    s.update(True)                      # second one halves
    assert s.scale == 8
    s.update(True)
    assert s.scale == 4
    for _ in range(2):
        s.update(False)
    assert s.scale == 4
    s.update(False)                     # third clean step since the last overflow: grow, hysteresis restored
    assert s.scale == 8 and s.cur_hysteresis == 2
    for _ in range(20):
        s.update(True)
    assert s.scale == 1                 # floor
    t = DynamicLossScaler(LossScalerConfig(initial_scale_power=4, scale_window=3))
    t.load_state_dict(s.state_dict())
    assert (t.scale, t.cur_iter, t.last_overflow_iter, t.skipped) == (s.scale, s.cur_iter, s.last_overflow_iter, s.skipped)


def test_scaler_factory_reads_ds_config_keys():
    assert make_loss_scaler(None) is None and make_loss_scaler({"enabled": False}) is None
    st = make_loss_scaler({"enabled": True, "loss_scale": 128})
    assert isinstance(st, LossScaler) and not st.dynamic and st.scale == 128
    dy = make_loss_scaler({"enabled": True, "loss_scale": 0, "initial_scale_power": 12, "loss_scale_window": 50,
                           "hysteresis": 1, "min_loss_scale": 2})
    assert dy.dynamic and dy.scale == 4096 and dy.cfg.scale_window == 50 and dy.cfg.min_scale == 2
    dy.update(True)
    assert dy.scale == 2048             # hysteresis 1: every overflow lowers the scale


def _engine(extra, ps=None, seed=5):
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=4, vocab_size=128, max_seq=32)
    ds = {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
          "optimizer": {"params": {"lr": 2e-3, "weight_decay": 0.0}}, "gradient_clipping": 1.0, **extra}
    return cfg, initialize(cfg, ds, ps, device=torch.device("cpu"), seed=seed)


def _batches(cfg, n=4, seed=0):
    g = torch.Generator().manual_seed(seed)
    mbs = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(n)]
    return [(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in mbs]


def test_fp16_run_matches_unscaled_run_and_skips_overflowing_steps(tmp_path):
    cfg, plain = _engine({})
    _, scaled = _engine({"fp16": {"enabled": True, "initial_scale_power": 10, "hysteresis": 1,
                                  "loss_scale_window": 2}})
    assert scaled.loss_scaler is not None and scaled.loss_scale == 1024
    it_a, it_b = itertools.cycle(_batches(cfg)), itertools.cycle(_batches(cfg))
    for _ in range(3):
        la, lb = float(plain.train_batch(it_a)), float(scaled.train_batch(it_b))
        assert abs(la - lb) < 1e-4, (la, lb)         # the reported loss is unscaled, updates identical
    assert scaled.loss_scale == 2048 and scaled.skipped_steps == 0   # window of 2 clean steps passed once
    # poison one gradient: the step is skipped, parameters and Adam state stay as they were
    before = scaled.flat.params.clone()
    steps_before = scaled.optimizer.step_count
    orig = scaled.flat.collect_autograd_grads

    def poisoned():
        orig()
        scaled.flat.grads[3] = float("inf")

    scaled.flat.collect_autograd_grads = poisoned
    scaled.train_batch(it_b)
    scaled.flat.collect_autograd_grads = orig
    assert scaled.skipped_steps == 1 and scaled.loss_scale == 1024
    assert torch.equal(before, scaled.flat.params) and scaled.optimizer.step_count == steps_before
    assert scaled.global_steps == 4
    # the scaler state travels with the optimizer state of a checkpoint
    scaled.save_checkpoint(str(tmp_path))
    _, fresh = _engine({"fp16": {"enabled": True, "initial_scale_power": 10, "hysteresis": 1,
                                 "loss_scale_window": 2}})
    fresh.load_checkpoint(str(tmp_path))
    assert fresh.loss_scale == 1024 and fresh.skipped_steps == 1
    assert fresh.loss_scaler.cur_iter == scaled.loss_scaler.cur_iter
    l1, l2 = float(scaled.train_batch(it_b)), float(fresh.train_batch(itertools.cycle(_batches(cfg))))
    assert abs(l1 - l2) < 1e-5


def _pp_eval_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    from b200flow.parallel.groups import ParallelState
    from test_parallel_cpu import _sync_weights_from_reference

    ps = ParallelState.initialize(tp=1, pp=2)
    cfg, eng = _engine({}, ps)
    _sync_weights_from_reference(eng, ps, cfg)
    it = itertools.cycle(_batches(cfg))
    ev = [float(eng.eval_batch(it, num_micro_batches=3)), float(eng.eval_batch(it))]
    tr = float(eng.train_batch(it))     # training after evaluation still works (p2p buffers reused)
    q.put((rank, ev, tr))
    dist.barrier()
    dist.destroy_process_group()


def test_eval_batch_under_pipeline_parallelism_matches_single_process():
    from test_parallel_cpu import _sync_weights_from_reference

    cfg, eng = _engine({})
    _sync_weights_from_reference(eng, None, cfg)
    it = itertools.cycle(_batches(cfg))
    want = [float(eng.eval_batch(it, num_micro_batches=3)), float(eng.eval_batch(it))]
    want_train = float(eng.train_batch(it))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_pp_eval_worker, args=(r, 2, 29641, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=300) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, ev, tr in got:            # every stage reports the same loss
        for a, b in zip(ev, want):
            assert abs(a - b) < 2e-3, (rank, ev, want)
        assert abs(tr - want_train) < 2e-3


def test_forward_backward_step_triple_equals_train_batch():
    """DeepSpeed's non-pipeline loop (`loss = engine(x, y); engine.backward(loss); engine.step()`) makes the
    same updates as `train_batch`, with `step()` acting only at accumulation boundaries."""
    cfg, a = _engine({"fp16": {"enabled": True, "initial_scale_power": 8}})
    _, b = _engine({"fp16": {"enabled": True, "initial_scale_power": 8}})
    data = _batches(cfg)
    it = itertools.cycle(data)
    la = [float(a.train_batch(it)) for _ in range(3)]            # gas = 2 -> 6 micro-batches
    lb, k = [], 0
    for _ in range(3):
        acc = 0.0
        for _ in range(2):
            tok, lab = data[k % len(data)]
            k += 1
            loss = b(tok, lab)
            b.backward(loss)
            assert b.is_gradient_accumulation_boundary() == (k % 2 == 0)
            b.step()                                              # no-op in the middle of an accumulation window
            acc += float(loss)
        lb.append(acc / 2)
    assert b.global_steps == 3 and a.global_steps == 3
    for x, y in zip(la, lb):
        assert abs(x - y) < 1e-5, (la, lb)
    assert torch.allclose(a.flat.params, b.flat.params, atol=1e-6)


def test_wall_clock_breakdown_and_flops_profiler_print(capsys):
    cfg, eng = _engine({"wall_clock_breakdown": True, "steps_per_print": 2,
                        "flops_profiler": {"enabled": True, "profile_step": 2}})
    it = itertools.cycle(_batches(cfg))
    for _ in range(4):
        eng.train_batch(it)
    out = capsys.readouterr().out
    lines = [l for l in out.splitlines() if l.startswith("time (ms) |")]
    assert len(lines) == 2 and all(k in lines[0] for k in ("fwd:", "bwd:", "step:"))
    assert out.count("flops profiler | step 2 |") == 1
    p = eng.flops_profile
    assert p["tokens_per_step"] == 2 * 32 * 2 and p["flops_per_step"] == cfg.flops_per_token(32) * p["tokens_per_step"]
    assert p["tflops_per_gpu"] > 0


def _comms_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    import contextlib
    import io

    from b200flow.parallel.groups import ParallelState

    ps = ParallelState.initialize(tp=2, pp=1)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        cfg, eng = _engine({"comms_logger": {"enabled": True}, "steps_per_print": 2}, ps)
        it = itertools.cycle(_batches(cfg))
        for _ in range(2):
            eng.train_batch(it)
        eng.comms_logger.disable()
    q.put((rank, buf.getvalue()))
    dist.barrier()
    dist.destroy_process_group()


def test_comms_logger_in_the_engine_reports_tensor_parallel_collectives():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_comms_worker, args=(r, 2, 29642, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert "comm op" in got[0] and "all_reduce" in got[0] and "all_gather" in got[0]   # SP all-gathers, norm all-reduce
    assert got[1].count("comm op") == 0                                                    # rank 0 prints


def test_activation_recompute_changes_memory_not_math():
    """`--checkpoint-activations` / `"activation_checkpointing": {"enabled": true}`: identical updates."""
    cfg, plain = _engine({})
    _, ckpt = _engine({"activation_checkpointing": {"enabled": True}})
    assert ckpt.model.recompute and not plain.model.recompute
    ia, ib = itertools.cycle(_batches(cfg)), itertools.cycle(_batches(cfg))
    for _ in range(3):
        la, lb = float(plain.train_batch(ia)), float(ckpt.train_batch(ib))
        assert abs(la - lb) < 1e-5, (la, lb)
    assert torch.allclose(plain.flat.params, ckpt.flat.params, atol=1e-6)


def test_accelerator_fp16_grad_scaler_skips_and_checkpoints(tmp_path):
    """Accelerate under `--mixed_precision fp16` (what the reference's BERT example passes): dynamic loss scale,
    inf / nan gradients skip optimizer and scheduler, `scaler.pt` is part of `save_state`."""
    from b200flow.engines.accelerate_api import Accelerator
    from b200flow.optim import AdamW, get_linear_schedule_with_warmup

    def build(mp_mode):
        torch.manual_seed(0)
        acc = Accelerator(cpu=True, mixed_precision=mp_mode)
        model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1))
        opt = AdamW(model.parameters(), lr=1e-2, weight_decay=0.0)
        sch = get_linear_schedule_with_warmup(opt, 2, 20)
        model, opt, sch = acc.prepare(model, opt, sch)
        return acc, model, opt, sch

    (a, ma, oa, sa), (b, mb, ob, sb) = build("no"), build("fp16")
    assert a.scaler is None and b.scaler.scale == 65536.0
    x, y = torch.randn(32, 8), torch.randn(32, 1)
    for _ in range(3):
        for acc, m, o, s in ((a, ma, oa, sa), (b, mb, ob, sb)):
            loss = torch.nn.functional.mse_loss(m(x), y)
            acc.backward(loss)
            acc.clip_grad_norm_(m.parameters(), 1.0)           # unscales first under fp16
            o.step()
            s.step()
            o.zero_grad()
    wa = torch.cat([p.detach().reshape(-1) for p in ma.parameters()])
    wb = torch.cat([p.detach().reshape(-1) for p in mb.parameters()])
    assert torch.allclose(wa, wb, atol=1e-5) and sa.get_last_lr() == sb.get_last_lr()
    # an overflowing step: parameters, Adam step count and the schedule stay put, the scale halves
    before, lr = wb.clone(), sb.get_last_lr()
    loss = torch.nn.functional.mse_loss(mb(x), y) * float("inf")
    b.backward(loss)
    ob.step()
    sb.step()
    ob.zero_grad()
    wb2 = torch.cat([p.detach().reshape(-1) for p in mb.parameters()])
    assert b.optimizer_step_was_skipped and torch.equal(before, wb2) and sb.get_last_lr() == lr
    assert b.scaler.scale == 32768.0
    out = b.save_state(str(tmp_path / "step_4"))
    assert (tmp_path / "step_4" / "scaler.pt").exists()
    c, mc, oc, sc = build("fp16")
    c.load_state(out)
    assert c.scaler.scale == 32768.0 and c.scaler.state_dict() == b.scaler.state_dict()


def test_megatron_lr_decay_styles():
    from b200flow.parallel.zero1 import cosine_lr, megatron_lr

    kw = dict(base_lr=1e-3, min_lr=1e-5, warmup_steps=10, decay_steps=110)
    for step in (1, 5, 10, 11, 60, 109, 110, 500):
        assert abs(megatron_lr(step, style="cosine", **kw) - cosine_lr(step, **kw)) < 1e-15
    assert megatron_lr(5, style="linear", **kw) == 5e-4                       # warm-up is shared
    assert abs(megatron_lr(60, style="linear", **kw) - (1e-5 + 0.5 * (1e-3 - 1e-5))) < 1e-12
    assert megatron_lr(60, style="constant", **kw) == 1e-3 and megatron_lr(10 ** 6, style="constant", **kw) == 1e-3
    assert abs(megatron_lr(40, style="inverse-square-root", **kw) - 1e-3 * (10 / 40) ** 0.5) < 1e-12
    assert megatron_lr(10 ** 9, style="inverse-square-root", **kw) == 1e-5
    assert megatron_lr(200, style="linear", **kw) == 1e-5


def test_fused_adamw_front_end_resumes_with_its_step_count():
    """`optimizer.bin` of an Accelerate checkpoint: the bias-correction step travels with the state, so a
    reloaded optimizer continues exactly (and its cached device tables are dropped)."""
    from b200flow.optim import AdamW

    def make():
        torch.manual_seed(0)
        w = torch.nn.Parameter(torch.randn(7, 5))
        return w, AdamW([w], lr=1e-2, weight_decay=0.01)

    def step(w, opt, k):
        w.grad = torch.full_like(w, 0.1 * (k + 1)) * w.detach().sign()
        opt.step()

    wa, oa = make()
    for k in range(3):
        step(wa, oa, k)
    import copy

    sd = copy.deepcopy(oa.state_dict())     # as if read back from optimizer.bin (state tensors are shared otherwise)
    assert sd["b2f_step"] == 3
    wb, ob = make()
    wb.data.copy_(wa.data)
    ob.load_state_dict(sd)
    assert ob._step == 3 and ob._tables == {}
    for k in range(3, 6):
        step(wa, oa, k)
        step(wb, ob, k)
    assert torch.allclose(wa, wb, atol=1e-7)


def _accum_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from b200flow.engines.accelerate_api import Accelerator
    from b200flow.optim import AdamW

    acc = Accelerator(cpu=True, gradient_accumulation_steps=2)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1))
    opt = AdamW(model.parameters(), lr=1e-2, weight_decay=0.0)
    model, opt = acc.prepare(model, opt)
    g = torch.Generator().manual_seed(1)
    x, y = torch.randn(2, 16, 8, generator=g), torch.randn(2, 16, 1, generator=g)       # 2 optimizer steps x 16 samples
    for step in range(2):
        for micro in range(2):                                                        # this rank's 2 micro-batches of 4
            lo = (rank * 2 + micro) * 4
            with acc.accumulate(model):
                loss = torch.nn.functional.mse_loss(model(x[step, lo:lo + 4]), y[step, lo:lo + 4])
                acc.backward(loss)
                opt.step()
                opt.zero_grad()
    if rank == 0:
        q.put(torch.cat([p.detach().reshape(-1) for p in model.parameters()]).tolist())
    acc.wait_for_everyone()
    dist.destroy_process_group()


def test_accumulation_over_ddp_equals_one_large_batch():
    """Accelerate's `accumulate` on two ranks (2 micro-batches each, reduction only on the last one) makes the
    update a single process makes on the 16-sample batch."""
    from b200flow.optim import AdamW

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1))
    opt = AdamW(model.parameters(), lr=1e-2, weight_decay=0.0)
    g = torch.Generator().manual_seed(1)
    x, y = torch.randn(2, 16, 8, generator=g), torch.randn(2, 16, 1, generator=g)
    for step in range(2):
        torch.nn.functional.mse_loss(model(x[step]), y[step]).backward()
        opt.step()
        opt.zero_grad()
    want = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_accum_worker, args=(r, 2, 29644, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = torch.tensor(q.get(timeout=300))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert torch.allclose(got, want, atol=1e-6), float((got - want).abs().max())


def test_hidden_dropout_masks_are_seeded_and_survive_recompute():
    """--hidden-dropout: training differs from the dropout-free run, evaluation does not; the masks are a function
    of the per-micro-batch seed, so activation recomputation reproduces them exactly."""
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    def engine(dropout, recompute=False):
        cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=4, vocab_size=128, max_seq=32,
                                 dropout=dropout)
        return cfg, initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
                                     "optimizer": {"params": {"lr": 2e-3}},
                                     "activation_checkpointing": {"enabled": recompute}},
                               device=torch.device("cpu"), seed=5)

    cfg, plain = engine(0.0)
    _, drop = engine(0.2)
    _, drop_ckpt = engine(0.2, recompute=True)
    data = _batches(cfg)
    e0 = float(plain.eval_batch(itertools.cycle(data)))
    assert abs(float(drop.eval_batch(itertools.cycle(data))) - e0) < 1e-6 and drop.model.training
    ia, ib, ic = (itertools.cycle(data) for _ in range(3))
    for step in range(3):
        la, lb, lc = float(plain.train_batch(ia)), float(drop.train_batch(ib)), float(drop_ckpt.train_batch(ic))
        assert abs(lb - lc) < 1e-6, (step, lb, lc)            # recompute draws the same masks
        assert abs(la - lb) > 1e-4                            # and dropout is really applied
    assert torch.allclose(drop.flat.params, drop_ckpt.flat.params, atol=1e-6)


def _dropout_tp_worker(rank, world, port, sp, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset
    from b200flow.parallel.groups import ParallelState

    ps = ParallelState.initialize(tp=2, pp=1)
    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=2, vocab_size=128, max_seq=32, dropout=0.3)
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2, "sequence_parallel": sp,
                           "optimizer": {"params": {"lr": 5e-3}}}, ps, device=torch.device("cpu"), seed=5)
    it = itertools.cycle(_batches(cfg))
    losses = [float(eng.train_batch(it)) for _ in range(3)]
    rep = torch.cat([p.detach().reshape(-1) for n, p in eng.model.named_parameters()
                     if n.startswith("embed.") or getattr(p, "tp_replicated", False)])
    both = [torch.empty_like(rep) for _ in range(2)]
    dist.all_gather(both, rep)
    q.put((rank, losses, bool(torch.equal(both[0], both[1])), float((rep - 1).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


def test_dropout_under_tensor_parallelism_keeps_replicated_state_identical():
    """Replicated activations (no sequence parallelism) need the same mask on every TP rank, sequence shards a
    different one per rank; either way the tensors every TP rank holds whole stay bit-identical while they train."""
    for sp, port in ((False, 29645), (True, 29646)):
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        procs = [ctx.Process(target=_dropout_tp_worker, args=(r, 2, port, sp, q)) for r in range(2)]
        for p in procs:
            p.start()
        res = [q.get(timeout=300) for _ in procs]
        for p in procs:
            p.join(timeout=120)
            assert p.exitcode == 0
        for rank, losses, same, moved in res:
            assert same, f"replicated tensors diverged between TP ranks (sequence_parallel={sp})"
            assert moved > 0 and all(v == v for v in losses)
        assert res[0][1] == res[1][1]                          # both ranks report the same loss


def _acc_api_worker(rank, world, port, tmp, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from b200flow.engines.accelerate_api import Accelerator

    acc = Accelerator(cpu=True)
    t = torch.arange(3 + rank, dtype=torch.float32)
    padded = acc.pad_across_processes(t, pad_index=-1)
    red = acc.reduce(torch.tensor([float(rank + 1)]), "mean")
    model = acc.prepare(torch.nn.Linear(4, 2))
    acc.save_model(model, tmp)
    ran = acc.on_main_process(lambda: "ran")()
    with acc.autocast():
        y = model(torch.ones(1, 4))
    q.put((rank, padded.tolist(), float(red), ran, acc.is_last_process, list(y.shape)))
    acc.wait_for_everyone()
    dist.destroy_process_group()


def test_accelerator_utility_calls(tmp_path):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_acc_api_worker, args=(r, 2, 29647, str(tmp_path), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res[0][1] == [0.0, 1.0, 2.0, -1.0] and res[1][1] == [0.0, 1.0, 2.0, 3.0]
    assert res[0][2] == res[1][2] == 1.5
    assert res[0][3] == "ran" and res[1][3] is None and res[1][4] and not res[0][4]
    assert (tmp_path / "model.safetensors").exists()


def test_generation_continues_a_memorised_sequence():
    """`GPTModel.generate` (Megatron's text-generation utilities at their simplest): a model trained on one periodic
    sequence continues it greedily; logits agree with the training loss; sampling respects top-k."""
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset

    cfg = config_from_preset("llama-tiny", hidden=64, heads=4, ffn=128, layers=2, vocab_size=32, max_seq=32)
    eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "optimizer": {"params": {"lr": 1e-2, "weight_decay": 0.0}}},
                     device=torch.device("cpu"), seed=1)
    period = torch.tensor([3, 9, 4, 17, 8, 21, 5, 11])
    seq = period.repeat(5)[:33]
    batch = torch.stack([seq, seq.roll(-2)[:33]])
    it = itertools.cycle([(batch[:, :-1].contiguous(), batch[:, 1:].contiguous())])
    for _ in range(120):
        loss = eng.train_batch(it)
    assert float(loss) < 0.05
    logits = eng.model.logits(batch[:, :-1])
    ce = torch.nn.functional.cross_entropy(logits.reshape(-1, 32).float(), batch[:, 1:].reshape(-1))
    assert abs(float(ce) - float(eng.eval_batch(it))) < 1e-4 and eng.model.training
    out = eng.model.generate(seq[None, :8], 12)
    assert out.shape == (1, 20) and out[0, 8:].tolist() == period.repeat(3)[8:20].tolist()
    sampled = eng.model.generate(seq[None, :8], 6, temperature=1.0, top_k=1, generator=torch.Generator().manual_seed(0))
    assert sampled[0].tolist() == out[0, :14].tolist()                      # top-1 sampling is greedy


def test_attention_dropout_mask_and_training_on_cpu():
    """--attention-dropout: the mask is a pure function of (seed, batch, head, query, key), drops the asked fraction,
    and a model trains with it; two runs with the same seeds give the same curve (what activation recomputation and a
    resumed run rely on)."""
    from b200flow.engines.ds_engine import initialize
    from b200flow.models.gpt import config_from_preset
    from b200flow.ops.attention import dropout_keep_mask

    keep, rinv = dropout_keep_mask(2, 3, 64, 0.1, 99)
    assert keep.shape == (2, 3, 64, 64) and abs((~keep).float().mean().item() - 26 / 256) < 0.02
    assert abs(rinv - 256 / 230) < 1e-9
    again, _ = dropout_keep_mask(2, 3, 64, 0.1, 99)
    other, _ = dropout_keep_mask(2, 3, 64, 0.1, 100)
    assert torch.equal(keep, again) and not torch.equal(keep, other)
    assert not torch.equal(keep[0, 0], keep[0, 1]) and not torch.equal(keep[0, 0], keep[1, 0])

    def curve():
        cfg = config_from_preset("gpt2-tiny")
        cfg.attention_dropout = 0.1
        eng = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "optimizer": {"params": {"lr": 2e-3}}},
                         device=torch.device("cpu"), seed=3)
        g = torch.Generator().manual_seed(5)
        toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(2)]
        it = itertools.cycle([(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks])
        return [float(eng.train_batch(it)) for _ in range(6)]

    a, b = curve(), curve()
    assert a == b and a[-1] < a[0]
    cfg = config_from_preset("gpt2-tiny")
    plain = initialize(cfg, {"train_micro_batch_size_per_gpu": 2, "optimizer": {"params": {"lr": 2e-3}}},
                       device=torch.device("cpu"), seed=3)
    g = torch.Generator().manual_seed(5)
    toks = [torch.randint(0, cfg.vocab_size, (2, 33), generator=g) for _ in range(2)]
    it = itertools.cycle([(t[:, :-1].contiguous(), t[:, 1:].contiguous()) for t in toks])
    assert abs(float(plain.train_batch(it)) - a[0]) > 1e-6          # the mask is really applied
