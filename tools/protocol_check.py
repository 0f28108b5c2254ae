"""Randomised interleaving check of the multi-rank Darlin protocol (csrc/cuda/darlin.cu:
`push_entries` / `pull_segment_replica` / `grid_wait_flow`): tagged (g, u) entries stored into the
landing table of every rank that holds or owns a column, double-buffered by step parity; every
rank sums the holders' entries itself and applies the same update to its replica; a writer starts
step s only after every peer finished step s-2 (the `done` flags).

It is a model of the protocol, not of the CUDA code: every rank is a little state machine whose
atomic actions (the pushes of a step, one entry poll, the end-of-step flag) are interleaved by a
seeded random scheduler; block orders deliberately repeat blocks in consecutive steps and two
steps apart, and ranks may hold no column of a block at all (so nothing but the flow control
keeps them from running ahead).  Checked on every schedule:
  * no deadlock: while some rank has work, some action is enabled;
  * an entry is never overwritten before every rank that needs it has read it;
  * a reader only ever accepts the entry of the step it is in, and sums exactly the holders'
    contributions of that step;
  * all replicas (and the owner's shard) of a coordinate hold the same value after every step.

    python tools/protocol_check.py [--ranks 3] [--keys 12] [--passes 4] [--schedules 300]
"""
from __future__ import annotations

import argparse
import random


class Model:
    def __init__(self, ranks: int, keys: int, blocks: int, passes: int, rng: random.Random, mutate: str = ""):
        self.R, self.K, self.rng, self.mutate = ranks, keys, rng, mutate
        self.owner = [k % ranks for k in range(keys)]
        # which workers hold nonzeros of a key (at least one); some ranks hold nothing of a block
        self.has = [set(r for r in range(ranks) if rng.random() < 0.6) or {rng.randrange(ranks)} for _ in range(keys)]
        self.block_of = [k % blocks for k in range(keys)]
        order = []
        for p in range(passes):
            b = list(range(blocks))
            rng.shuffle(b)
            if p == 0 and blocks >= 2:  # prior-group style repeats: a, a, b, a
                b = [b[0], b[0], b[1], b[0]] + b
            order.append(b)
        self.passes = order
        # landing[q][parity][r][k] = (tag, value) as rank q sees it; readers[...] = ranks that still
        # have to read the current content
        self.land = [[[{} for _ in range(ranks)] for _ in range(2)] for _ in range(ranks)]
        self.unread = [[[{} for _ in range(ranks)] for _ in range(2)] for _ in range(ranks)]
        self.done = [[0] * ranks for _ in range(ranks)]      # done[q][r]: rank q's view of rank r
        self.replica = [[0] * keys for _ in range(ranks)]    # every rank's copy of the coordinate
        self.shard = [0] * keys                               # the owner's authoritative copy
        self.contrib = {}                                     # (step, r, k) -> value pushed
        self.pc = [self._program(r) for r in range(ranks)]
        self.pos = [0] * ranks
        self.errors = []

    def needers(self, k):
        return self.has[k] | {self.owner[k]}

    # ---- per-rank programs: (kind, seq, key) with seq = 1, 2, ... over all passes ----------------
    def _program(self, r: int):
        prog, seq = [], 0
        for p, blocks in enumerate(self.passes):
            for b in blocks:
                seq += 1
                keys = [k for k in range(self.K) if self.block_of[k] == b]
                prog.append(("flow", seq, None))            # wait: every peer finished seq - 2
                prog.append(("push", seq, [k for k in keys if r in self.has[k]]))
                for k in keys:
                    if r in self.needers(k):
                        prog.append(("pull", seq, k))
                prog.append(("done", seq, None))
            prog.append(("barrier", seq, p))                # rank barrier at the end of a pass
        return prog

    def enabled(self, r: int) -> bool:
        kind, seq, arg = self.pc[r][self.pos[r]]
        if kind == "flow":
            if self.mutate == "no_flow":
                return True
            return all(self.done[r][q] + 2 >= seq for q in range(self.R))
        if kind == "pull":
            par = seq & 1 if self.mutate != "no_parity" else 0
            return all(self.land[r][par][h].get(arg, (0, None))[0] >= seq for h in self.has[arg])
        if kind == "barrier":
            return all(self.pos[q] >= len(self.pc[q]) or self.pc[q][self.pos[q]][1] >= seq for q in range(self.R))
        return True

    def step(self, r: int) -> None:
        kind, seq, arg = self.pc[r][self.pos[r]]
        if kind == "push":
            par = seq & 1 if self.mutate != "no_parity" else 0
            for k in arg:
                v = self.rng.randrange(1, 1000)
                self.contrib[(seq, r, k)] = v
                for q in self.needers(k):
                    if self.unread[q][par][r].get(k):
                        self.errors.append(f"seq {seq}: rank {r} overwrites key {k} on rank {q} before it was read "
                                           f"(tag {self.land[q][par][r][k][0]})")
                    self.land[q][par][r][k] = (seq, v)
                    self.unread[q][par][r][k] = True
        elif kind == "pull":
            par = seq & 1 if self.mutate != "no_parity" else 0
            total = 0
            for h in sorted(self.has[arg]):
                tag, v = self.land[r][par][h][arg]
                if tag != seq:
                    self.errors.append(f"seq {seq}: rank {r} reads key {arg} of worker {h} with tag {tag}")
                total += v
                self.unread[r][par][h][arg] = False
            want = sum(self.contrib[(seq, h, arg)] for h in self.has[arg])
            if total != want:
                self.errors.append(f"seq {seq}: rank {r} summed {total} for key {arg}, the step pushed {want}")
            self.replica[r][arg] += total                     # the deterministic update
            if self.owner[arg] == r:
                self.shard[arg] += total
        elif kind == "done":
            for q in range(self.R):
                self.done[q][r] = seq
        elif kind == "barrier":
            pass
        self.pos[r] += 1

    def run(self):
        while True:
            live = [r for r in range(self.R) if self.pos[r] < len(self.pc[r])]
            if not live:
                break
            ready = [r for r in live if self.enabled(r)]
            if not ready:
                self.errors.append("deadlock: " + str({r: self.pc[r][self.pos[r]][:2] for r in live}))
                break
            # rank 0 is a straggler: picked ten times less often, so that the others run ahead of it
            self.step(self.rng.choices(ready, weights=[0.1 if r == 0 else 1.0 for r in ready])[0])
            if self.errors:
                break
        if not self.errors:
            for k in range(self.K):
                vals = {self.replica[r][k] for r in self.needers(k)} | {self.shard[k]}
                if len(vals) != 1:
                    self.errors.append(f"end: copies of key {k} differ: {sorted(vals)}")
        return self.errors


def check(ranks=3, keys=12, blocks=4, passes=4, schedules=300, seed=0, mutate="", verbose=True) -> int:
    bad = 0
    for i in range(schedules):
        rng = random.Random(seed * 100003 + i)
        errs = Model(ranks, keys, blocks, passes, rng=rng, mutate=mutate).run()
        if errs:
            bad += 1
            if verbose:
                print(f"schedule {i}: {errs[0]}")
    return bad


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=3)
    ap.add_argument("--keys", type=int, default=12)
    ap.add_argument("--blocks", type=int, default=4)
    ap.add_argument("--passes", type=int, default=4)
    ap.add_argument("--schedules", type=int, default=300)
    ap.add_argument("--mutate", default="", choices=["", "no_flow", "no_parity"],
                    help="break the protocol on purpose: the checker must notice")
    a = ap.parse_args()
    n = check(a.ranks, a.keys, a.blocks, a.passes, a.schedules, mutate=a.mutate)
    print("violations:", n)
    raise SystemExit(1 if n else 0)
