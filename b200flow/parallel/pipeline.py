"""Pipeline parallelism: 1F1B schedule over torch.distributed p2p (NCCL on GPUs).

The reference gets this from DeepSpeed's `PipelineEngine.train_batch()` +
`TrainSchedule` + `pipe/p2p.py` when `--pipeline-model-parallel-size 2` is set
(examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml:40; SURVEY.md
§2F "PP", §3.1).  By design (BASELINE.json north star) stage-boundary traffic stays
on NCCL send/recv; what this module adds is the instruction stream and the
deadlock-free pairing of sends and receives (`batch_isend_irecv` groups each
steady-state send with the receive that the peer is blocked on).

`schedule_1f1b` is a pure function returning the instruction list of one stage, so
the schedule itself is unit-tested without any process group.

Bubble filling (`schedule_zb`).  A backward pass is two halves per linear: dgrad, which the
previous stage is waiting for, and wgrad, which only the optimizer needs.  With the halves split
("B" = dgrad only, "W" = the queued wgrad GEMMs of one micro-batch, `ops.gemm.defer_wgrad`) the
1F1B order is kept, stage s runs W_j right after B_{j+s}, and whatever is left runs at the end.
The last stage then hands the final gradients upstream before doing its last weight gradients,
and earlier stages fill their wait for the next gradient with weight gradients: the idle time per
stage falls from (p-1)(f + b) to (p-1)(f + dgrad - wgrad), i.e. to about a third for
f = dgrad = wgrad (the "ZB-H1" schedule of the zero-bubble pipeline papers); `simulate` is the
event model the unit tests hold both schedules against.

Why the exchanges stay rendezvous-at-need on the compute stream.  NCCL send / recv are kernels
that occupy SMs until the peer shows up.  The GEMMs here are persistent kernels with one CTA (pair)
per SM and a static tile loop, so a receive posted far ahead of its data, or a send posted far ahead
of its receive, parks a communication kernel on an SM for tens of milliseconds and every GEMM in that
window loses the SM to a straggler CTA.  Posting each send together with the receive the stage needs
next keeps those kernels short-lived (both sides arrive within microseconds of each other), and the
time a stage spends inside them is time it had to wait for its peer's result anyway: per micro-batch
the exposed cost is one 16 MiB NVLink transfer (tens of microseconds) against ~100 ms of compute.
What the schedule can win is the bubble, hence `schedule_zb`.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class Instr:
    op: str   # "F" | "B"
    mb: int


def schedule_1f1b(stage: int, stages: int, micro_batches: int) -> list[Instr]:
    """Instruction stream of `stage`: warm-up forwards, 1F1B steady state, cool-down backwards."""
    warm = min(stages - stage - 1, micro_batches)
    out = [Instr("F", i) for i in range(warm)]
    for i in range(micro_batches - warm):
        out.append(Instr("F", warm + i))
        out.append(Instr("B", i))
    out += [Instr("B", micro_batches - warm + i) for i in range(warm)]
    return out


def schedule_zb(stage: int, stages: int, micro_batches: int) -> list[Instr]:
    """1F1B order with split backward: "B" is dgrad only, "W" the deferred weight gradients of a micro-batch;
    stage s runs W_j right after B_{j+s} and the remainder at the end."""
    out, nb, nw = [], 0, 0
    for ins in schedule_1f1b(stage, stages, micro_batches):
        out.append(ins)
        if ins.op == "B":
            nb += 1
            if nb - nw > stage:
                out.append(Instr("W", nw))
                nw += 1
    out += [Instr("W", j) for j in range(nw, micro_batches)]
    return out


def bubble_fraction(stages: int, micro_batches: int) -> float:
    return (stages - 1) / (micro_batches + stages - 1)


def simulate(stages: int, micro_batches: int, schedule, f: float = 1.0, dgrad: float = 1.0, wgrad: float = 1.0) -> float:
    """Makespan of one batch under the dependency model: F_i on stage s after F_i on s-1; B_i on s after F_i on s and
    B_i on s+1; W_i after B_i.  A schedule without "W" instructions pays dgrad + wgrad inside "B"."""
    lists = [schedule(s, stages, micro_batches) for s in range(stages)]
    split = any(i.op == "W" for l in lists for i in l)
    done: dict[tuple, float] = {}
    t, idx = [0.0] * stages, [0] * stages
    moved = True
    while moved:
        moved = False
        for s in range(stages):
            while idx[s] < len(lists[s]):
                ins = lists[s][idx[s]]
                deps = []
                if ins.op == "F" and s > 0:
                    deps.append(("F", s - 1, ins.mb))
                if ins.op == "B":
                    deps.append(("F", s, ins.mb))
                    if s < stages - 1:
                        deps.append(("B", s + 1, ins.mb))
                if ins.op == "W":
                    deps.append(("B", s, ins.mb))
                if any(d not in done for d in deps):
                    break
                start = max([t[s]] + [done[d] for d in deps])
                dur = {"F": f, "B": dgrad + (0.0 if split else wgrad), "W": wgrad}[ins.op]
                t[s] = done[(ins.op, s, ins.mb)] = start + dur
                idx[s] += 1
                moved = True
    assert all(idx[s] == len(lists[s]) for s in range(stages)), "schedule deadlocks"
    return max(t)


class P2P:
    """Stage-boundary communication with fixed activation shape."""

    def __init__(self, prev_rank: int | None, next_rank: int | None, shape, dtype, device, group=None):
        self.prev, self.next = prev_rank, next_rank
        self.shape, self.dtype, self.device, self.group = tuple(shape), dtype, device, group

        self._connected = False

    def connect(self):
        """Open both directions of every stage boundary before the schedule starts, one direction per call.  NCCL
        connects a peer lazily inside the first group that names it; with a single micro-batch the first stage's first
        group is "send forward + receive backward" while its neighbour's is "receive forward" alone, and that
        asymmetric first contact fails inside NCCL's bootstrap exchange (2.28: "Message truncated").  After this the
        schedule may mix directions freely."""
        if self._connected or not dist.is_initialized():
            return
        self._connected = True
        one = torch.zeros(1, dtype=self.dtype, device=self.device)
        ops = []
        if self.next is not None:
            ops.append(dist.P2POp(dist.isend, one, self.next, self.group))
        if self.prev is not None:
            ops.append(dist.P2POp(dist.irecv, torch.empty_like(one), self.prev, self.group))
        self._run(ops)
        ops = []
        if self.prev is not None:
            ops.append(dist.P2POp(dist.isend, one, self.prev, self.group))
        if self.next is not None:
            ops.append(dist.P2POp(dist.irecv, torch.empty_like(one), self.next, self.group))
        self._run(ops)

    def _buf(self):
        return torch.empty(self.shape, dtype=self.dtype, device=self.device)

    def _run(self, ops):
        if not ops:
            return
        for r in dist.batch_isend_irecv(ops):
            r.wait()

    def exchange(self, send_fwd=None, send_bwd=None, recv_fwd=False, recv_bwd=False):
        """Post every requested send/recv as ONE group; returns (activation_in, grad_in)."""
        self.connect()
        ops, act_in, grad_in = [], None, None
        for t in (send_fwd, send_bwd):
            if t is not None and (tuple(t.shape) != self.shape or t.dtype != self.dtype):
                raise ValueError(f"pipeline boundary tensor {tuple(t.shape)} {t.dtype} does not match the agreed "
                                 f"{self.shape} {self.dtype}")
        if send_fwd is not None and self.next is not None:
            ops.append(dist.P2POp(dist.isend, send_fwd.contiguous(), self.next, self.group))
        if send_bwd is not None and self.prev is not None:
            ops.append(dist.P2POp(dist.isend, send_bwd.contiguous(), self.prev, self.group))
        if recv_fwd and self.prev is not None:
            act_in = self._buf()
            ops.append(dist.P2POp(dist.irecv, act_in, self.prev, self.group))
        if recv_bwd and self.next is not None:
            grad_in = self._buf()
            ops.append(dist.P2POp(dist.irecv, grad_in, self.next, self.group))
        self._run(ops)
        return act_in, grad_in


class _Trace:
    """B2F_PP_TRACE=<prefix>: device timestamps of every schedule instruction ("x" = the exchange that follows an
    F / B), one JSON line per batch and stage in <prefix>.stage<k>.jsonl - the 2-GPU timeline evidence."""

    def __init__(self, stage):
        self.stage, self.ev = stage, []
        self.mark("start", -1)

    def mark(self, op, mb):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        self.ev.append((op, mb, e))

    def dump(self, prefix):
        import json

        self.mark("end", -1)
        torch.cuda.synchronize()
        t0 = self.ev[0][2]
        rows = [(op, mb, round(t0.elapsed_time(e), 3)) for op, mb, e in self.ev]
        with open(f"{prefix}.stage{self.stage}.jsonl", "a") as f:
            f.write(json.dumps(rows) + "\n")


def run_1f1b(stage: int, stages: int, micro_batches: int, forward_fn, p2p: P2P, zero_bubble: bool = False):
    """Execute one training batch on this stage.

    forward_fn(mb_index, activation_in_or_None) -> output tensor (activation for the next
    stage, or the scalar loss on the last stage).  Returns the list of last-stage losses
    (empty on other stages).  Gradients accumulate into the parameters' main_grad.
    `zero_bubble`: run `schedule_zb` - backward passes queue their weight-gradient GEMMs and the
    "W" instructions launch them later.
    """
    import os

    from ..ops import gemm as G

    first, last = stage == 0, stage == stages - 1
    instrs = (schedule_zb if zero_bubble else schedule_1f1b)(stage, stages, micro_batches)
    trace = _Trace(stage) if (os.environ.get("B2F_PP_TRACE") and p2p.device.type == "cuda") else None
    inputs: dict[int, torch.Tensor] = {}
    outputs: dict[int, torch.Tensor] = {}
    wqueues: dict[int, list] = {}
    losses = []
    pending_act = None      # activation received ahead of its forward
    pending_grad = None     # gradient received ahead of its backward
    for k, ins in enumerate(instrs):
        # the exchange after this instruction is paired with what the next F / B needs ("W" has no traffic)
        nxt = next((i for i in instrs[k + 1:] if i.op != "W"), None)
        if ins.op == "W":
            if ins.mb in wqueues:         # (not already run ahead of a receive-only exchange, see below)
                trace and trace.mark("W", ins.mb)
                G.run_deferred(wqueues.pop(ins.mb))
            continue
        trace and trace.mark(ins.op, ins.mb)
        if ins.op == "F":
            if not first:
                act = pending_act
                pending_act = None
                if act is None:
                    act, _ = p2p.exchange(recv_fwd=True)
                act.requires_grad_(True)
            else:
                act = None
            out = forward_fn(ins.mb, act)
            inputs[ins.mb], outputs[ins.mb] = act, out
            if last:
                losses.append(out.detach())
            trace and trace.mark("x", ins.mb)
            # send the activation on; pair it with the receive the next instruction needs
            want_grad = (not last) and nxt is not None and nxt.op == "B"
            want_act = (not first) and nxt is not None and nxt.op == "F"
            a, g = p2p.exchange(send_fwd=None if last else out.detach(), recv_bwd=want_grad,
                                recv_fwd=want_act)
            if a is not None:
                pending_act = a
            if g is not None:
                pending_grad = g
        else:
            out = outputs.pop(ins.mb)
            act = inputs.pop(ins.mb)
            grad = None
            if not last:
                grad = pending_grad
                pending_grad = None
                if grad is None:
                    _, grad = p2p.exchange(recv_bwd=True)
            if zero_bubble:
                with G.defer_wgrad(wqueues.setdefault(ins.mb, [])):
                    out.backward() if last else torch.autograd.backward(out, grad)
            elif last:
                out.backward()
            else:
                torch.autograd.backward(out, grad)
            dact = act.grad if act is not None else None
            if dact is None or first:
                # nothing to send upstream: the exchange below only receives, i.e. only waits.  Weight gradients that
                # the schedule placed right behind this backward run first, so the wait is spent computing (the first
                # stage's cool-down would otherwise idle in the receive with its queue full).
                for w in instrs[k + 1:]:
                    if w.op != "W":
                        break
                    if w.mb in wqueues:
                        trace and trace.mark("W", w.mb)
                        G.run_deferred(wqueues.pop(w.mb))
            trace and trace.mark("x", ins.mb)
            want_act = (not first) and nxt is not None and nxt.op == "F"
            want_grad = (not last) and nxt is not None and nxt.op == "B"
            a, g = p2p.exchange(send_bwd=dact, recv_fwd=want_act, recv_bwd=want_grad)
            if a is not None:
                pending_act = a
            if g is not None:
                pending_grad = g
    assert not wqueues, "weight gradients left in the queue"
    trace and trace.dump(os.environ["B2F_PP_TRACE"])
    return losses


def run_forward_only(stage: int, stages: int, micro_batches: int, forward_fn, p2p: P2P):
    """Evaluation schedule (DeepSpeed `InferenceSchedule` in spirit): every micro-batch flows
    through the stages once, no backward and nothing kept.  Returns the last stage's losses."""
    first, last = stage == 0, stage == stages - 1
    losses = []
    with torch.no_grad():
        for mb in range(micro_batches):
            act = None
            if not first:
                act, _ = p2p.exchange(recv_fwd=True)
            out = forward_fn(mb, act)
            if last:
                losses.append(out.detach())
            else:
                p2p.exchange(send_fwd=out.detach())
    return losses
