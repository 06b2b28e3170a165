"""Explicit-state models of the buffer-reuse protocols of the peer-memory kernels (K1, K2, K3, K4, K5), explored over EVERY
interleaving of the ranks' steps.

The kernels reuse a fixed number of symmetric buffers across calls and never reset a flag: flags carry the call counter
(epoch), buffers alternate with its parity.  DESIGN.md argues why two buffers suffice; this file states each protocol as a
small program per rank and checks the argument mechanically: whatever the scheduling of ranks (and of the steps that run
concurrently inside one kernel), every read returns the data of the call it belongs to, and no rank waits forever.  The
same models with ONE buffer are expected to fail - that is the evidence that the check can fail at all.

What the models assume, as the kernels do: steps of one rank follow stream order; a flag or counter update is visible
after the data it covers (release / acquire); a wait is a spin on `value >= epoch`.
The GPU-side counterpart is the race-delay build (tools/race_check.sh).
usage: python tools/protocol_model.py
"""
from __future__ import annotations


class Violation(Exception):
    pass


def explore(initial, successors, is_final, limit: int = 2_000_000):
    """Depth-first search of the reachable states.  `successors(state)` yields next states (raising Violation on a bad
    read); a non-final state without successors is a deadlock.  Returns the number of states visited."""
    seen, stack = {initial}, [initial]
    while stack:
        s = stack.pop()
        nxt = list(successors(s))
        if not nxt and not is_final(s):
            raise Violation(f"deadlock in state {s}")
        for n in nxt:
            if n not in seen:
                if len(seen) >= limit:
                    raise RuntimeError("state space larger than the limit")
                seen.add(n)
                stack.append(n)
    return len(seen)


def _set(tup, i, v):
    return tup[:i] + (v,) + tup[i + 1:]


# ---------------------------------------------------------------------------------------------- K1: in-place all-reduce
def k1_model(world: int = 2, calls: int = 3, end_barrier: bool = True):
    """Per rank and call e (csrc/comm/collectives.cu::allreduce_f32_kernel, one block): the gradients of the call are in
    the rank's symmetric buffer (one sub-slice per rank); barrier with flag value 2e-1 ("all gradients final"); the rank
    reads sub-slice `me` of every rank, writes the average into sub-slice `me` of every rank; barrier with 2e ("all
    slices landed"); then it reads its whole buffer.  One flag word per (rank, peer), never reset.
    `end_barrier=False` is the negative control."""
    W = world
    # steps: 0 produce, 1 announce A, 2..W+1 wait A (each peer incl. self), W+2 reduce (read all), W+3 write all,
    #        W+4 announce B, W+5..2W+4 wait B, 2W+5 consume
    per_call = 2 * W + 6

    def initial():
        pcs = tuple((1, 0) for _ in range(W))
        buf = tuple(tuple(None for _ in range(W)) for _ in range(W))         # buf[rank][sub-slice]
        flag = tuple(tuple(0 for _ in range(W)) for _ in range(W))           # flag[at][from]
        return (pcs, buf, flag)

    def successors(state):
        pcs, buf, flag = state
        for r in range(W):
            e, k = pcs[r]
            if e > calls:
                continue
            nxt = (e, k + 1) if k + 1 < per_call else (e + 1, 0)
            if k == 0:
                yield (_set(pcs, r, nxt), _set(buf, r, tuple(("g", r, e) for _ in range(W))), flag)
            elif k == 1 or k == W + 4:
                val = 2 * e - 1 if k == 1 else 2 * e
                yield (_set(pcs, r, nxt), buf, tuple(_set(flag[at], r, val) for at in range(W)))
            elif 2 <= k <= W + 1:
                if flag[r][k - 2] >= 2 * e - 1:
                    yield (_set(pcs, r, nxt), buf, flag)
            elif k == W + 2:
                for p in range(W):
                    if buf[p][r] != ("g", p, e):
                        raise Violation(f"K1: rank {r} call {e} reduced {buf[p][r]} from rank {p}")
                yield (_set(pcs, r, nxt), buf, flag)
            elif k == W + 3:
                yield (_set(pcs, r, nxt), tuple(_set(buf[p], r, ("avg", r, e)) for p in range(W)), flag)
            elif W + 5 <= k <= 2 * W + 4:
                if not end_barrier or flag[r][k - W - 5] >= 2 * e:
                    yield (_set(pcs, r, nxt), buf, flag)
            else:
                for q in range(W):
                    if buf[r][q] != ("avg", q, e):
                        raise Violation(f"K1: rank {r} call {e} consumed {buf[r][q]} in sub-slice {q}")
                yield (_set(pcs, r, nxt), buf, flag)

    return initial(), successors, lambda s: all(e > calls for e, _ in s[0])


# ---------------------------------------------------------------------------------------------- K2: ZeRO-1 step
def k2_model(world: int = 2, calls: int = 3, start_barrier: bool = True):
    """Per rank and step e (csrc/comm/collectives.cu zero1_reduce / zero1_adam_gather, parallel/fused_dp.py):
         forward + backward: read my copy of every parameter shard (must be the values of step e-1), write my gradients
         barrier 2e-1 ("all gradients final" - also: everybody is done reading its parameters)
         reduce: read gradient shard `me` from every rank
         Adam on shard `me`, store the updated shard into every rank's parameter buffer
         barrier 2e ("all shards landed" - also: everybody is done reading my gradients).
       `start_barrier=False` is the negative control (a fast rank overwrites parameters a slow rank is still using)."""
    W = world
    per_call = 2 * W + 6        # 0 compute, 1 announce A, 2..W+1 wait A, W+2 reduce, W+3 write params, W+4 announce B, W+5..2W+4 wait B

    def initial():
        pcs = tuple((1, 0) for _ in range(W))
        params = tuple(tuple(("p", q, 0) for q in range(W)) for _ in range(W))     # params[rank][shard]
        grads = tuple(tuple(None for _ in range(W)) for _ in range(W))             # grads[rank][shard]
        flag = tuple(tuple(0 for _ in range(W)) for _ in range(W))
        return (pcs, params, grads, flag)

    def successors(state):
        pcs, params, grads, flag = state
        for r in range(W):
            e, k = pcs[r]
            if e > calls:
                continue
            nxt = (e, k + 1) if k + 1 < per_call - 1 else (e + 1, 0)
            if k == 0:
                for q in range(W):
                    if params[r][q] != ("p", q, e - 1):
                        raise Violation(f"K2: rank {r} step {e} computed with parameter shard {params[r][q]}")
                yield (_set(pcs, r, nxt), params, _set(grads, r, tuple(("g", r, e) for _ in range(W))), flag)
            elif k == 1 or k == W + 4:
                val = 2 * e - 1 if k == 1 else 2 * e
                yield (_set(pcs, r, nxt), params, grads, tuple(_set(flag[at], r, val) for at in range(W)))
            elif 2 <= k <= W + 1:
                if not start_barrier or flag[r][k - 2] >= 2 * e - 1:
                    yield (_set(pcs, r, nxt), params, grads, flag)
            elif k == W + 2:
                for p in range(W):
                    if grads[p][r] != ("g", p, e):
                        raise Violation(f"K2: rank {r} step {e} reduced gradient {grads[p][r]} of rank {p}")
                yield (_set(pcs, r, nxt), params, grads, flag)
            elif k == W + 3:
                yield (_set(pcs, r, nxt), tuple(_set(params[p], r, ("p", r, e)) for p in range(W)), grads, flag)
            else:
                if flag[r][k - W - 5] >= 2 * e:
                    yield (_set(pcs, r, nxt), params, grads, flag)

    return initial(), successors, lambda s: all(e > calls for e, _ in s[0])


# ---------------------------------------------------------------------------------------------- K3: all-gather -> GEMM
def k3_model(world: int = 2, calls: int = 3, slots: int = 2):
    """Per rank and call e: copy the shard into slot[e % slots]; announce ready = e to everybody; then for every peer:
    wait for its announcement of e and pull its slot[e % slots] (csrc/comm/fused_gemm.cu::ag_comm_role,
    parallel/fused_tp.py::ag_gemm).  State: per rank (call, step), slot contents, ready flags."""
    peers = [[p for p in range(world) if p != r] for r in range(world)]
    per_call = 2 + 2 * (world - 1)          # copy, announce, (wait, read) per peer

    def initial():
        pcs = tuple((1, 0) for _ in range(world))
        slot = tuple(tuple(None for _ in range(slots)) for _ in range(world))
        ready = tuple(tuple(0 for _ in range(world)) for _ in range(world))      # ready[at][from]
        return (pcs, slot, ready)

    def successors(state):
        pcs, slot, ready = state
        for r in range(world):
            e, k = pcs[r]
            if e > calls:
                continue
            nxt_pc = (e, k + 1) if k + 1 < per_call else (e + 1, 0)
            if k == 0:
                yield (_set(pcs, r, nxt_pc), _set(slot, r, _set(slot[r], e % slots, (r, e))), ready)
            elif k == 1:
                rd = tuple(_set(ready[at], r, e) for at in range(world))
                yield (_set(pcs, r, nxt_pc), slot, rd)
            else:
                p = peers[r][(k - 2) // 2]
                if (k - 2) % 2 == 0:
                    if ready[r][p] >= e:
                        yield (_set(pcs, r, nxt_pc), slot, ready)
                else:
                    if slot[p][e % slots] != (p, e):
                        raise Violation(f"K3: rank {r} call {e} pulled {slot[p][e % slots]} from rank {p}")
                    yield (_set(pcs, r, nxt_pc), slot, ready)

    return initial(), successors, lambda s: all(e > calls for e, _ in s[0])


# ---------------------------------------------------------------------------------------------- K4: GEMM -> reduce-scatter
def k4_model(world: int = 2, calls: int = 3, buffers: int = 2):
    """Per rank and call e: for every peer, store the partial rows into the peer's receive buffer [e % buffers][me] and
    then its flag = e; afterwards, for every peer, wait for flag[e % buffers][peer] >= e and read the partial
    (csrc/comm/fused_gemm.cu RSPolicy, parallel/fused_tp.py::gemm_rs)."""
    peers = [[p for p in range(world) if p != r] for r in range(world)]
    n = world - 1
    per_call = 4 * n

    def initial():
        pcs = tuple((1, 0) for _ in range(world))
        recv = tuple(tuple(tuple(None for _ in range(world)) for _ in range(buffers)) for _ in range(world))
        flag = tuple(tuple(tuple(0 for _ in range(world)) for _ in range(buffers)) for _ in range(world))
        return (pcs, recv, flag)

    def successors(state):
        pcs, recv, flag = state
        for r in range(world):
            e, k = pcs[r]
            if e > calls:
                continue
            b = e % buffers
            nxt_pc = (e, k + 1) if k + 1 < per_call else (e + 1, 0)
            if k < 2 * n:
                p = peers[r][k // 2]
                if k % 2 == 0:
                    yield (_set(pcs, r, nxt_pc), _set(recv, p, _set(recv[p], b, _set(recv[p][b], r, (r, e)))), flag)
                else:
                    yield (_set(pcs, r, nxt_pc), recv, _set(flag, p, _set(flag[p], b, _set(flag[p][b], r, e))))
            else:
                p = peers[r][(k - 2 * n) // 2]
                if (k - 2 * n) % 2 == 0:
                    if flag[r][b][p] >= e:
                        yield (_set(pcs, r, nxt_pc), recv, flag)
                else:
                    if recv[r][b][p] != (p, e):
                        raise Violation(f"K4: rank {r} call {e} read {recv[r][b][p]} from rank {p}")
                    yield (_set(pcs, r, nxt_pc), recv, flag)

    return initial(), successors, lambda s: all(e > calls for e, _ in s[0])


# ---------------------------------------------------------------------------------------------- K5: fused MoE
def k5_model(world: int = 2, calls: int = 3, buffers: int = 2, epoch_flags: bool = True):
    """Per rank and call e (parallel/fused_moe.py::FusedMoE.run, csrc/comm/fused_moe.cu):
         Z   clear the completion counter of the NEXT call's parity
         UP  kernel: dispatch steps - for every owner o: store tokens into o's token buffer [e % buffers][me], then the
             unit flag (epoch valued, one region) = e - run concurrently with the expert-GEMM steps - for every
             source s: wait flag[s] >= e, read tokens [e % buffers][s]
         DOWN kernel (after UP completed): for every source s: store result rows into s's home buffer
             [e % buffers][me], then add 1 to s's completion counter [e % buffers]
         W   wait for `world` arrivals on my counter [e % buffers], read every home row.
       A rank sends to itself as well (its own experts), as the kernels do.
       `epoch_flags=False` is the negative control: a boolean flag that is never reset lets an owner read the tokens of
       the previous call.  (With ONE buffer the model still passes: a rank's rows only come home after every owner has
       consumed its tokens, so the round trip is a barrier per call; the implementation keeps two parities anyway and
       does not lean on that.)"""
    W = world

    def initial():
        # per rank: (call, stage, dispatch_done bitmask, gemm_done bitmask, down_idx)
        pcs = tuple((1, 0, 0, 0, 0) for _ in range(W))
        tok = tuple(tuple(tuple(None for _ in range(W)) for _ in range(buffers)) for _ in range(W))
        flag = tuple(tuple(0 for _ in range(W)) for _ in range(W))                # flag[owner][source]
        rows = tuple(tuple(tuple(None for _ in range(W)) for _ in range(buffers)) for _ in range(W))
        done = tuple(tuple(0 for _ in range(buffers)) for _ in range(W))
        return (pcs, tok, flag, rows, done)

    full = (1 << W) - 1

    def successors(state):
        pcs, tok, flag, rows, done = state
        for r in range(W):
            e, stage, dmask, gmask, di = pcs[r]
            if e > calls:
                continue
            b = e % buffers
            if stage == 0:                                   # Z
                nb = (e + 1) % buffers
                nd = _set(done, r, _set(done[r], nb, 0)) if buffers > 1 else done
                yield (_set(pcs, r, (e, 1, 0, 0, 0)), tok, flag, rows, nd)
            elif stage == 1:                                 # UP kernel: dispatch and GEMM steps in any order
                for o in range(W):
                    bit = 1 << o
                    if not dmask & bit:                      # store tokens + flag to owner o (two steps folded: data
                        t2 = _set(tok, o, _set(tok[o], b, _set(tok[o][b], r, (r, e))))     # is visible before the flag)
                        f2 = _set(flag, o, _set(flag[o], r, e if epoch_flags else 1))
                        nm = dmask | bit
                        st = 2 if (nm == full and gmask == full) else 1
                        yield (_set(pcs, r, (e, st, nm, gmask, 0)), t2, f2, rows, done)
                for s in range(W):
                    bit = 1 << s
                    if not gmask & bit and flag[r][s] >= (e if epoch_flags else 1):
                        if tok[r][b][s] != (s, e):
                            raise Violation(f"K5: owner {r} call {e} read tokens {tok[r][b][s]} of source {s}")
                        nm = gmask | bit
                        st = 2 if (dmask == full and nm == full) else 1
                        yield (_set(pcs, r, (e, st, dmask, nm, 0)), tok, flag, rows, done)
            elif stage == 2:                                 # DOWN kernel: rows home, one source after the other
                s = di
                r2 = _set(rows, s, _set(rows[s], b, _set(rows[s][b], r, (r, s, e))))
                d2 = _set(done, s, _set(done[s], b, done[s][b] + 1))
                nxt = (e, 2, dmask, gmask, di + 1) if di + 1 < W else (e, 3, dmask, gmask, 0)
                yield (_set(pcs, r, nxt), tok, flag, r2, d2)
            else:                                            # W: all my rows are home
                if done[r][b] >= W:
                    for o in range(W):
                        if rows[r][b][o] != (o, r, e):
                            raise Violation(f"K5: source {r} call {e} read home row {rows[r][b][o]} of owner {o}")
                    nd = done if buffers > 1 else _set(done, r, _set(done[r], b, 0))
                    yield (_set(pcs, r, (e + 1, 0, 0, 0, 0)), tok, flag, rows, nd)

    return initial(), successors, lambda s: all(p[0] > calls for p in s[0])


def main():
    for name, make in (("K1", k1_model), ("K2", k2_model), ("K3", k3_model), ("K4", k4_model), ("K5", k5_model)):
        for world, calls in ((2, 4), (3, 3)):
            n = explore(*make(world, calls))
            print(f"{name}: world {world}, {calls} calls, as built: {n} states, every read correct, no deadlock")
        try:
            kw = {"K1": {"end_barrier": False}, "K2": {"start_barrier": False}, "K3": {"slots": 1}, "K4": {"buffers": 1}, "K5": {"epoch_flags": False}}[name]
            explore(*make(2, 3, **kw))
            print(f"{name}: negative control {kw} passed (unexpected)")
        except Violation as v:
            print(f"{name}: negative control {kw} fails as it must: {v}")


if __name__ == "__main__":
    main()
