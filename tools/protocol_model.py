#!/usr/bin/env python3
"""Exhaustive interleaving check of the cross-rank protocol of the sharded data-parallel step.

compute-sanitizer's racecheck cannot look across processes (and is not available on the development pool), so the
ordering argument of `csrc/p2p_allreduce.cu` / `parallel/p2p.py` / `controller.py` is checked on a model instead: every
rank is two sequential threads (training stream, reducer side stream) of atomic actions; the checker walks EVERY
interleaving of all ranks' actions (depth-first, memoised on the full state) and fails on

  * a read that finds a cell in another state than the protocol promises — at the start OR at the end of the read
    (reads are split in two actions, so a write that lands while a read is in flight is seen),
  * a write into a cell that somebody may still need (wrong predecessor state),
  * a state without an enabled action before every rank has finished (deadlock).

What one step of rank r does (W ranks, B buckets, slice s of a bucket is owned by rank s):

  training stream                                     side stream
  ---------------                                     -----------
  per bucket b (last first):
    forward / backward of b reads parameters P[r][b][*]
    and writes the gradients G[r][b][*]      --event-->  announce flag0[b] to the peers
                                                          wait for the peers' flag0[b]
                                                          read G[p][b][r] of every rank p, write the sum to G[r][b][r]
  wait for the side stream
  statistics tail: announce, wait, read every rank's tail, signal "done reading", wait for the peers' "done", write own tail
  sharded Adam: read G[r][b][r], write the new P[p][b][r] into EVERY rank p, clear G[r][b][*]
  completion round: signal, wait for every peer's signal          (the next step's forward may now read P[r])

`pipelined=True` models `--pipelined-update`: the sharded Adam of step t is issued bucket by bucket on the side stream at
the start of step t+1 (its scalars were parked in separate device words, so the tail may be cleared), the forward pass
waits per bucket ("need") for the kernel that delivers that bucket's parameters, and the last parked update is flushed on
the training stream at the end.

`drop=` removes one element of the protocol; every one of them must make the checker fail (tests/test_cpu_core.py),
which shows both that the model can see such bugs and that none of the synchronisation is decorative.
"""
from __future__ import annotations

import sys
from typing import List, Optional, Tuple

ZERO, GRAD, RED = 0, 1, 2          # state of a gradient cell (tagged with the step that produced it)
VAL = 1                            # statistics tail: local value / reduced value use GRAD(=VAL) / RED


class Model:
    def __init__(self, W: int, B: int, T: int, drop: Optional[str] = None, pipelined: bool = False):
        self.W, self.B, self.T, self.drop, self.pipelined = W, B, T, drop, pipelined
        self.main_prog = [self._main(r) for r in range(W)]
        self.side_prog = [self._side(r) for r in range(W)]

    # ---- programs ----------------------------------------------------------------------------------------------
    def _adam(self, t, buckets, slot) -> List[Tuple]:
        prog = [("adam_begin", t, buckets), ("adam_write", t, buckets)]
        if self.drop != "adam_round":
            prog += [("adam_sig", t, slot), ("adam_wait", t, slot)]
        return prog

    def _main(self, r) -> List[Tuple]:
        prog = []
        allb = tuple(range(self.B))
        for t in range(self.T):
            if self.pipelined:
                prog.append(("apply_event", t))          # the side stream may start the parked update of step t-1
            prog.append(("tail_zero", t))
            for b in reversed(range(self.B)):
                if self.pipelined and t > 0 and self.drop != "need":
                    prog.append(("need", t, b))
                prog += [("bwd_begin", t, b), ("bwd_end", t, b)]
            if self.drop != "join":
                prog.append(("join", t))
            prog += [("tail_ann", t)]
            if self.drop != "tail_wait":
                prog.append(("tail_wait", t))
            prog += [("tail_read_begin", t), ("tail_read_end", t)]
            if self.drop != "tail_done":
                prog += [("tail_done_sig", t), ("tail_done_wait", t)]
            prog += [("tail_write", t)]
            if self.pipelined:
                prog.append(("park", t))
            else:
                prog += self._adam(t, allb, self.B)
        if self.pipelined:                                 # flush_pending_update(): on the training stream
            prog += self._adam(self.T - 1, allb, self.B)
        prog.append(("final_read", self.T))
        return prog

    def _side(self, r) -> List[Tuple]:
        prog = []
        self.side_join = []                                # side_join[t]: program counter after step t's last reduce-scatter
        for t in range(self.T):
            if self.pipelined and t > 0:
                prog.append(("wait_apply_event", t))
                for b in reversed(range(self.B)):
                    prog += self._adam(t - 1, (b,), b)
                    prog.append(("adam_done", t, b))
            for b in reversed(range(self.B)):
                prog.append(("rs_ann", t, b))
                if self.drop != "rs_wait":
                    prog.append(("rs_wait", t, b))
                prog += [("rs_read_begin", t, b), ("rs_read_end", t, b)]
            self.side_join.append(len(prog))
        return prog

    # ---- state -------------------------------------------------------------------------------------------------
    # pcs: (main pc, side pc) per rank; ev[r] = number of buckets whose backward has finished (over all steps);
    # flags: flag0[r][b][src], tail_a[r][src], tail_d[r][src], adam_f[r][src] hold epochs; G[r][b][s] = (kind, step);
    # P[r][b][s] = version; Tv[r] = (kind, step)
    def initial(self):
        W, B = self.W, self.B
        pcs = tuple((0, 0) for _ in range(W))
        ev = tuple(0 for _ in range(W))
        flag0 = tuple(tuple(tuple(0 for _ in range(W)) for _ in range(B)) for _ in range(W))
        fl = tuple(tuple(0 for _ in range(W)) for _ in range(W))
        G = tuple(tuple(tuple((ZERO, -1) for _ in range(W)) for _ in range(B)) for _ in range(W))
        P = tuple(tuple(tuple(0 for _ in range(W)) for _ in range(B)) for _ in range(W))
        Tv = tuple((ZERO, -1) for _ in range(W))
        adam_f = tuple(tuple(tuple(0 for _ in range(W)) for _ in range(B + 1)) for _ in range(W))     # [rank][slot][src]
        aux = tuple((0, -1) + tuple(0 for _ in range(B)) for _ in range(W))      # (apply events, parked step, done[b]...)
        return (pcs, ev, flag0, fl, fl, adam_f, G, P, Tv, aux)

    @staticmethod
    def _set2(t, i, v):
        return t[:i] + (v,) + t[i + 1:]

    def _set3(self, t, i, j, v):
        return self._set2(t, i, self._set2(t[i], j, v))

    def _set4(self, t, i, j, k, v):
        return self._set2(t, i, self._set3(t[i], j, k, v))

    # returns ("blocked", None) | ("ok", new_state) | ("violation", message)
    def step(self, st, r, which):
        pcs, ev, flag0, tail_a, tail_d, adam_f, G, P, Tv, aux = st
        W, B = self.W, self.B
        pc_main, pc_side = pcs[r]
        prog = self.main_prog[r] if which == 0 else self.side_prog[r]
        pc = pc_main if which == 0 else pc_side
        if pc >= len(prog):
            return "blocked", None
        act = prog[pc]
        name, t = act[0], act[1]
        e = t + 1
        peers = [p for p in range(W) if p != r]

        def bad(msg):
            return "violation", "rank %d %s: %s" % (r, act, msg)

        if name == "tail_zero":
            Tv = self._set2(Tv, r, (ZERO, t))
        elif name in ("bwd_begin", "bwd_end", "final_read"):
            bs = range(B) if name == "final_read" else [act[2]]
            for b in bs:
                if any(P[r][b][s] != t for s in range(W)):
                    return bad("parameters of bucket %d are %s, expected version %d everywhere" % (b, P[r][b], t))
            if name == "bwd_end":
                b = act[2]
                for s in range(W):
                    if G[r][b][s][0] != ZERO:
                        return bad("gradient cell [%d][%d] not cleared (%s) when backward accumulates" % (b, s, G[r][b][s]))
                    G = self._set4(G, r, b, s, (GRAD, t))
                ev = self._set2(ev, r, ev[r] + 1)
        elif name == "join":
            if pc_side < self.side_join[t]:                   # wait_stream(side)
                return "blocked", None
        elif name == "apply_event":
            aux = self._set2(aux, r, (aux[r][0] + 1,) + aux[r][1:])
        elif name == "wait_apply_event":
            if aux[r][0] < t + 1:
                return "blocked", None
        elif name == "park":
            if Tv[r] != (RED, t):
                return bad("parks a norm from a tail in state %s" % (Tv[r],))
            aux = self._set2(aux, r, (aux[r][0], t) + aux[r][2:])
        elif name == "adam_done":
            b = act[2]
            aux = self._set2(aux, r, aux[r][:2 + b] + (t,) + aux[r][3 + b:])
        elif name == "need":
            if aux[r][2 + act[2]] < t:
                return "blocked", None
        elif name == "rs_ann":
            b = act[2]
            need = t * B + (B - 1 - b) + 1                  # the event of this bucket's backward
            if ev[r] < need:
                return "blocked", None
            for p in peers:
                flag0 = self._set4(flag0, p, b, r, e)
        elif name == "rs_wait":
            b = act[2]
            if any(flag0[r][b][p] < e for p in peers):
                return "blocked", None
        elif name in ("rs_read_begin", "rs_read_end"):
            b = act[2]
            for p in range(W):
                if G[p][b][r] != (GRAD, t):
                    return bad("reads rank %d's gradients of bucket %d in state %s, expected (GRAD, %d)" % (p, b, G[p][b][r], t))
            if name == "rs_read_end":
                G = self._set4(G, r, b, r, (RED, t))
        elif name == "tail_ann":
            Tv = self._set2(Tv, r, (VAL, t)) if Tv[r] == (ZERO, t) else Tv
            if Tv[r] != (VAL, t):
                return bad("statistics tail in state %s" % (Tv[r],))
            for p in peers:
                tail_a = self._set3(tail_a, p, r, e)
        elif name == "tail_wait":
            if any(tail_a[r][p] < e for p in peers):
                return "blocked", None
        elif name in ("tail_read_begin", "tail_read_end"):
            for p in range(W):
                if Tv[p] != (VAL, t):
                    return bad("reads rank %d's statistics in state %s, expected (VAL, %d)" % (p, Tv[p], t))
        elif name == "tail_done_sig":
            for p in peers:
                tail_d = self._set3(tail_d, p, r, e)
        elif name == "tail_done_wait":
            if any(tail_d[r][p] < e for p in peers):
                return "blocked", None
        elif name == "tail_write":
            Tv = self._set2(Tv, r, (RED, t))
        elif name == "adam_begin":
            if self.pipelined:
                if aux[r][1] != t:
                    return bad("applies an update whose scalars were parked for step %d" % aux[r][1])
            elif Tv[r] != (RED, t):
                return bad("norm read from a tail in state %s" % (Tv[r],))
            for b in act[2]:
                if G[r][b][r] != (RED, t):
                    return bad("own slice of bucket %d is %s, expected (RED, %d)" % (b, G[r][b][r], t))
        elif name == "adam_write":
            for b in act[2]:
                for p in range(W):
                    if P[p][b][r] != t:
                        return bad("parameter slice [%d][%d][%d] is version %d, expected %d" % (p, b, r, P[p][b][r], t))
                    P = self._set4(P, p, b, r, t + 1)
                for s in range(W):
                    G = self._set4(G, r, b, s, (ZERO, t))
        elif name == "adam_sig":
            for p in peers:
                adam_f = self._set4(adam_f, p, act[2], r, e)
        elif name == "adam_wait":
            if any(adam_f[r][act[2]][p] < e for p in peers):
                return "blocked", None
        else:
            raise AssertionError(name)
        pcs = self._set2(pcs, r, (pc_main + 1, pc_side) if which == 0 else (pc_main, pc_side + 1))
        return "ok", (pcs, ev, flag0, tail_a, tail_d, adam_f, G, P, Tv, aux)

    def finished(self, st):
        return all(st[0][r] == (len(self.main_prog[r]), len(self.side_prog[r])) for r in range(self.W))


def check(W: int = 2, B: int = 2, T: int = 2, drop: Optional[str] = None, pipelined: bool = False,
          max_states: int = 5_000_000):
    """Returns (violation message or None, number of distinct states visited)."""
    m = Model(W, B, T, drop, pipelined)
    init = m.initial()
    seen = {init}
    stack = [init]
    while stack:
        st = stack.pop()
        progressed = False
        for r in range(W):
            for which in (0, 1):
                kind, out = m.step(st, r, which)
                if kind == "violation":
                    return out, len(seen)
                if kind == "ok":
                    progressed = True
                    if out not in seen:
                        seen.add(out)
                        stack.append(out)
        if not progressed and not m.finished(st):
            return "deadlock with program counters %s" % (st[0],), len(seen)
        if len(seen) > max_states:
            raise RuntimeError("state space larger than %d" % max_states)
    return None, len(seen)



# ---- inside one kernel: the one-shot all-reduce and the number of co-resident CTAs ------------------------------------
def check_one_shot(W: int = 2, grid: int = 3, resident: int = 3, epochs: int = 2):
    """Block-level model of the one-shot kernel (`p2p_allreduce_kernel`, one_shot branch): every CTA waits for the peers'
    announcement, reads, bumps the per-bucket counter; the CTA that finds the counter at grid-1 resets it, advances the
    epoch word and signals "done reading"; EVERY CTA then waits for the peers' "done" before it overwrites its part.
    A CTA occupies one of `resident` slots from its first action until it retires.  Returns (violation or None, states).

    With grid <= resident there is no deadlock in any interleaving; with grid > resident the resident CTAs of every rank
    wait for a signal that only a not-yet-resident CTA of a peer can cause — the reason `b200_p2p_allreduce` refuses
    one-shot launches of more than kOneShotMaxBlocks CTAs."""
    # per rank: epoch word, counter, per-CTA pc (0 = not started, 1 = waiting flag0, 2 = read, 3 = counted, 4 = waiting
    # flag1, 5 = retired); launches are sequential per rank (stream order)
    init = (tuple(0 for _ in range(W)),                         # completed launches per rank
            tuple(tuple(0 for _ in range(grid)) for _ in range(W)),   # CTA pcs of the running launch
            tuple(0 for _ in range(W)),                         # counter
            tuple(0 for _ in range(W)),                         # epoch word
            tuple(tuple(0 for _ in range(W)) for _ in range(W)),      # flag0[r][src]
            tuple(tuple(0 for _ in range(W)) for _ in range(W)),      # flag1[r][src]
            tuple(tuple(0 for _ in range(grid)) for _ in range(W)))   # epoch value each CTA read at its start
    seen, stack = {init}, [init]

    def set2(t, i, v):
        return t[:i] + (v,) + t[i + 1:]

    while stack:
        st = stack.pop()
        done, pcs, counter, epoch_word, flag0, flag1, myep = st
        if all(d == epochs for d in done):
            continue
        progressed = False
        for r in range(W):
            if done[r] == epochs:
                continue
            peers = [p for p in range(W) if p != r]
            live = sum(1 for pc in pcs[r] if 0 < pc < 5)
            for c in range(grid):
                pc = pcs[r][c]
                nxt = None
                if pc == 0 and live < resident and (c == 0 or pcs[r][c - 1] > 0):      # CTAs start in index order
                    e = epoch_word[r] + 1
                    f0 = flag0
                    if c == 0:                                                        # block 0 announces
                        for p in peers:
                            f0 = set2(f0, p, set2(f0[p], r, e))
                    nxt = (done, set2(pcs, r, set2(pcs[r], c, 1)), counter, epoch_word, f0, flag1,
                           set2(myep, r, set2(myep[r], c, e)))
                elif pc == 1 and all(flag0[r][p] >= myep[r][c] for p in peers):
                    nxt = (done, set2(pcs, r, set2(pcs[r], c, 2)), counter, epoch_word, flag0, flag1, myep)
                elif pc == 2:                                                          # read done -> count
                    last = counter[r] == grid - 1
                    cnt = 0 if last else counter[r] + 1
                    ew, f1 = epoch_word, flag1
                    if last:
                        if myep[r][c] != epoch_word[r] + 1:
                            return "rank %d: CTAs of one launch disagree about the epoch" % r, len(seen)
                        ew = set2(epoch_word, r, myep[r][c])
                        for p in peers:
                            f1 = set2(f1, p, set2(f1[p], r, myep[r][c]))
                    nxt = (done, set2(pcs, r, set2(pcs[r], c, 4)), set2(counter, r, cnt), ew, flag0, f1, myep)
                elif pc == 4 and all(flag1[r][p] >= myep[r][c] for p in peers):
                    npcs = set2(pcs[r], c, 5)
                    nd = done
                    if all(x == 5 for x in npcs):                                      # launch retired: next one may start
                        nd = set2(done, r, done[r] + 1)
                        npcs = tuple(0 for _ in range(grid))
                    nxt = (nd, set2(pcs, r, npcs), counter, epoch_word, flag0, flag1, myep)
                if nxt is not None:
                    progressed = True
                    if nxt not in seen:
                        seen.add(nxt)
                        stack.append(nxt)
        if not progressed:
            return "deadlock: CTA states %s" % (pcs,), len(seen)
    return None, len(seen)


DROPS = ("join", "rs_wait", "tail_wait", "tail_done", "adam_round")
DROPS_PIPELINED = DROPS + ("need",)


def main():
    W = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    T = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    for pipelined in (False, True):
        what = "pipelined update" if pipelined else "update at the end of the step"
        v, n = check(W, B, T, pipelined=pipelined)
        print("%s, %d ranks x %d buckets x %d steps: %s (%d states)" % (what, W, B, T, v or "no violation in any interleaving", n))
        for d in (DROPS_PIPELINED if pipelined else DROPS):
            v, n = check(W, B, T, drop=d, pipelined=pipelined)
            print("  without %-11s: %s (%d states)" % (d, v or "NO VIOLATION FOUND", n))
    for grid, resident in ((3, 3), (4, 3)):
        v, n = check_one_shot(W, grid, resident)
        print("one-shot kernel, %d CTAs, %d co-resident: %s (%d states)" % (grid, resident, v or "no deadlock in any interleaving", n))


if __name__ == "__main__":
    main()
