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


def bubble_fraction(stages: int, micro_batches: int) -> float:
    return (stages - 1) / (micro_batches + stages - 1)


class P2P:
    """Stage-boundary communication with fixed activation shape."""

    def __init__(self, prev_rank: int | None, next_rank: int | None, shape, dtype, device, group=None):
        self.prev, self.next = prev_rank, next_rank
        self.shape, self.dtype, self.device, self.group = tuple(shape), dtype, device, group

    def _buf(self):
        return torch.empty(self.shape, dtype=self.dtype, device=self.device)

    def _run(self, ops):
        if not ops:
            return
        for r in dist.batch_isend_irecv(ops):
            r.wait()

    def exchange(self, send_fwd=None, send_bwd=None, recv_fwd=False, recv_bwd=False):
        """Post every requested send/recv as ONE group; returns (activation_in, grad_in)."""
        ops, act_in, grad_in = [], None, None
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


def run_1f1b(stage: int, stages: int, micro_batches: int, forward_fn, p2p: P2P):
    """Execute one training batch on this stage.

    forward_fn(mb_index, activation_in_or_None) -> output tensor (activation for the next
    stage, or the scalar loss on the last stage).  Returns the list of last-stage losses
    (empty on other stages).  Gradients accumulate into the parameters' main_grad.
    """
    first, last = stage == 0, stage == stages - 1
    instrs = schedule_1f1b(stage, stages, micro_batches)
    inputs: dict[int, torch.Tensor] = {}
    outputs: dict[int, torch.Tensor] = {}
    losses = []
    pending_act = None      # activation received ahead of its forward
    pending_grad = None     # gradient received ahead of its backward
    for k, ins in enumerate(instrs):
        nxt = instrs[k + 1] if k + 1 < len(instrs) else None
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
            if last:
                out.backward()
            else:
                grad = pending_grad
                pending_grad = None
                if grad is None:
                    _, grad = p2p.exchange(recv_bwd=True)
                torch.autograd.backward(out, grad)
            dact = act.grad if act is not None else None
            want_act = (not first) and nxt is not None and nxt.op == "F"
            want_grad = (not last) and nxt is not None and nxt.op == "B"
            a, g = p2p.exchange(send_bwd=dact, recv_fwd=want_act, recv_bwd=want_grad)
            if a is not None:
                pending_act = a
            if g is not None:
                pending_grad = g
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
