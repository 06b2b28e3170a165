"""Deadlock-freedom of the pipeline's p2p exchanges for a grid of (stages, micro-batches, schedule): `run_1f1b` is run
per stage against a recording stand-in for `P2P` (no communication), which yields each stage's sequence of exchange
groups exactly as the engine would post them; the groups are then matched under the strictest semantics - every send
and receive is a rendezvous, a stage sits in ONE group at a time (the engine waits on a group before it goes on), the
i-th send A->B pairs with the i-th receive of B from A - and every stage must drain its sequence."""
import pytest
import torch

from b200flow.ops import gemm as G
from b200flow.parallel import pipeline as PP


class RecordingP2P:
    def __init__(self, stage, stages, log):
        self.stage, self.stages, self.log = stage, stages, log
        self.prev = stage - 1 if stage > 0 else None
        self.next = stage + 1 if stage < stages - 1 else None

    def exchange(self, send_fwd=None, send_bwd=None, recv_fwd=False, recv_bwd=False):
        ops, act, grad = [], None, None
        if send_fwd is not None and self.next is not None:
            ops.append(("send", self.next))
        if send_bwd is not None and self.prev is not None:
            ops.append(("send", self.prev))
        if recv_fwd and self.prev is not None:
            ops.append(("recv", self.prev))
            act = torch.zeros(2, 3)
        if recv_bwd and self.next is not None:
            ops.append(("recv", self.next))
            grad = torch.zeros(2, 3)
        if ops:
            self.log.append(ops)
        return act, grad


def exchange_groups(stage, stages, mbs, zero_bubble):
    log = []
    w = torch.nn.Parameter(torch.ones(3))
    last = stage == stages - 1

    def forward(i, act):
        x = act if act is not None else torch.ones(2, 3)
        y = x * w
        return y.sum() if last else y

    old = G.DEFER_CAPABLE
    try:
        PP.run_1f1b(stage, stages, mbs, forward, RecordingP2P(stage, stages, log), zero_bubble=zero_bubble)
    finally:
        G.DEFER_CAPABLE = old
    return log


def drain(per_stage):
    """Rendezvous matching of the recorded groups.  Returns the stages that still hold unmatched groups (empty: ok)."""
    pos = [0] * len(per_stage)
    done = [set() for _ in per_stage]            # indices of matched ops inside the current group of each stage
    sent = {}                                    # (a, b) -> number of matched a->b transfers (pairs match in order)
    progress = True
    while progress:
        progress = False
        for a, groups in enumerate(per_stage):
            if pos[a] >= len(groups):
                continue
            for i, (kind, b) in enumerate(groups[pos[a]]):
                if i in done[a] or kind != "send" or pos[b] >= len(per_stage[b]):
                    continue
                # the earliest unmatched receive-from-a in b's CURRENT group
                for j, (k2, src) in enumerate(per_stage[b][pos[b]]):
                    if k2 == "recv" and src == a and j not in done[b]:
                        # in-order pairing: no earlier unmatched send a->b in a's current group
                        if any(g == ("send", b) and q not in done[a] for q, g in enumerate(groups[pos[a]][:i])):
                            break
                        done[a].add(i)
                        done[b].add(j)
                        sent[(a, b)] = sent.get((a, b), 0) + 1
                        progress = True
                        break
            for s in range(len(per_stage)):
                if pos[s] < len(per_stage[s]) and len(done[s]) == len(per_stage[s][pos[s]]):
                    pos[s] += 1
                    done[s] = set()
                    progress = True
    return [s for s in range(len(per_stage)) if pos[s] < len(per_stage[s])], sent


@pytest.mark.parametrize("zero_bubble", [False, True])
@pytest.mark.parametrize("stages", [2, 3, 4])
def test_exchanges_never_deadlock(stages, zero_bubble):
    for mbs in (1, 2, 3, 4, 5, 8, 16):
        groups = [exchange_groups(s, stages, mbs, zero_bubble) for s in range(stages)]
        stuck, sent = drain(groups)
        assert not stuck, f"stages {stuck} stuck with {stages} stages, {mbs} micro-batches, zero_bubble={zero_bubble}"
        for a in range(stages - 1):              # every activation went forward and every gradient came back
            assert sent[(a, a + 1)] == mbs and sent[(a + 1, a)] == mbs


def test_the_matcher_reports_a_deadlock():
    # both sides receive first and only then send: a rendezvous deadlock
    stuck, _ = drain([[[("recv", 1)], [("send", 1)]], [[("recv", 0)], [("send", 0)]]])
    assert stuck == [0, 1]
