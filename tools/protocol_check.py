"""Randomised interleaving check of the multi-rank Darlin protocol (csrc/cuda/darlin.cu,
`server_phase_direct` / `mirror_store` / `mirror_poll`): landing slots that are their own arrival
flags, consumer reset, tagged mirror entries, active bits kept in step on both sides.

It is a model of the protocol, not of the CUDA code: every rank is a little state machine whose
atomic actions (one push, one slot poll+reset, one publish, one mirror poll) are interleaved by a
seeded random scheduler.  Checked on every schedule:
  * no deadlock: while some rank has work, some action is enabled;
  * a push never lands on a slot that still holds an unconsumed value;
  * the owner sums exactly the contributions of the current visit of the key;
  * a worker never accepts a mirror entry of another step;
  * worker and owner copies of the active bit agree whenever a step starts.

    python tools/protocol_check.py [--ranks 3] [--keys 12] [--passes 4] [--schedules 300]
"""
from __future__ import annotations

import argparse
import random

EMPTY = None


class Model:
    def __init__(self, ranks: int, keys: int, blocks: int, passes: int, revive: bool, rng: random.Random,
                 mutate: str = ""):
        self.R, self.K, self.rng, self.mutate = ranks, keys, rng, mutate
        self.owner = [k % ranks for k in range(keys)]
        # which workers hold nonzeros of a key (at least one)
        self.has = [set(r for r in range(ranks) if rng.random() < 0.7) or {rng.randrange(ranks)} for _ in range(keys)]
        self.block_of = [k % blocks for k in range(keys)]
        order = []
        for p in range(passes):
            b = list(range(blocks))
            rng.shuffle(b)
            order += [(p, x) for x in b]
        self.steps = order                      # global step s -> (pass, block)
        self.reset_pass = {p for p in range(1, passes) if rng.random() < 0.5}   # KKT filter resets
        self.revive = revive
        self.land = {(self.owner[k], r, k): EMPTY for k in range(keys) for r in self.has[k]}
        self.mirror = {(r, k): (0, None) for k in range(keys) for r in self.has[k]}   # (tag, value)
        self.w_active = [[True] * keys for _ in range(ranks)]      # worker copies
        self.s_active = [True] * keys                              # owner copy
        self.s_marked = [False] * keys                             # weight is the NaN mark
        self.filter_at = {k: rng.randrange(len(order)) for k in range(keys) if rng.random() < 0.4}
        self.pc = [self._program(r) for r in range(ranks)]
        self.pos = [0] * ranks
        self.errors = []

    # ---- per-rank programs: lists of (kind, step, key, extra) ------------------------------------
    def _program(self, r: int):
        prog = []
        for s, (p, b) in enumerate(self.steps):
            first_of_pass = s == 0 or self.steps[s - 1][0] != p
            if first_of_pass and p in self.reset_pass:
                prog.append(("reset", s, None, None))
            keys = [k for k in range(self.K) if self.block_of[k] == b]
            prog.append(("begin", s, None, None))
            for k in keys:
                if r in self.has[k]:
                    prog.append(("push", s, k, None))
            for k in keys:
                if self.owner[k] == r:
                    for w in sorted(self.has[k]):
                        prog.append(("poll", s, k, w))
                    prog.append(("publish", s, k, None))
            for k in keys:
                if r in self.has[k]:
                    prog.append(("pull", s, k, None))
            prog.append(("barrier", s, None, None))
        return prog

    # ---- one atomic action; returns False when it is not enabled yet ------------------------------
    def try_step(self, r: int) -> bool:
        kind, s, k, x = self.pc[r][self.pos[r]]
        tag = s + 1
        if kind == "reset":                      # pass begin: refill the active sets (both copies live on r)
            for kk in range(self.K):
                self.w_active[r][kk] = True
                if self.owner[kk] == r:
                    self.s_active[kk] = True
                    if self.revive:
                        self.s_marked[kk] = False
        elif kind == "begin":
            for kk in range(self.K):
                if self.owner[kk] == r and self.block_of[kk] == self.steps[s][1]:
                    self._acc = getattr(self, "_acc", {})
                    self._acc[(kk, s)] = []
        elif kind == "push":
            if not self.w_active[r][k]:
                pass                              # nothing is sent for an inactive column
            else:
                slot = (self.owner[k], r, k)
                if self.land[slot] is not EMPTY:
                    self.errors.append(f"step {s}: push of rank {r} key {k} lands on an unconsumed value")
                self.land[slot] = (s, r)
        elif kind == "poll":
            if not self.s_active[k]:
                pass                              # owner's bit clear: nobody pushes, nothing to poll
            else:
                slot = (r, x, k)
                if self.land[slot] is EMPTY:
                    return False                  # spin
                if self.land[slot] != (s, x):
                    self.errors.append(f"step {s}: owner {r} reads {self.land[slot]} for key {k} worker {x}")
                self._acc[(k, s)].append(x)
                if self.mutate != "no_reset":
                    self.land[slot] = EMPTY       # consumer reset
        elif kind == "publish":
            if self.s_active[k]:
                if sorted(self._acc[(k, s)]) != sorted(self.has[k]):
                    self.errors.append(f"step {s}: key {k} summed {self._acc[(k, s)]}, holders {self.has[k]}")
                republish = self.s_marked[k] and self.mutate != "no_republish"
                if self.s_marked[k] and not republish:
                    self.s_active[k] = False      # the bug the kernel comment warns about: workers that
                    self.pos[r] += 1              # pushed after a reset would wait for ever
                    return True
                if self.s_marked[k] or self.filter_at.get(k) == s:
                    self.s_active[k] = False      # (still) filtered: mark goes out again
                    self.s_marked[k] = True
                    val = "mark"
                else:
                    val = ("w", s)
                for w in self.has[k]:
                    self.mirror[(w, k)] = (tag, val)
        elif kind == "pull":
            if self.w_active[r][k]:
                t, val = self.mirror[(r, k)]
                if t != tag:
                    if t > tag:
                        self.errors.append(f"step {s}: rank {r} sees a future entry for key {k}")
                    return False                  # spin on the tag
                if val == "mark":
                    self.w_active[r][k] = False
                elif val != ("w", s):
                    self.errors.append(f"step {s}: rank {r} accepted {val} for key {k}")
        elif kind == "barrier":
            pass                                  # local barrier C: nothing to model across ranks
        self.pos[r] += 1
        return True

    def run(self) -> list:
        while True:
            live = [r for r in range(self.R) if self.pos[r] < len(self.pc[r])]
            if not live:
                break
            self.rng.shuffle(live)
            for r in live:
                if self.try_step(r):
                    break
            else:
                where = {r: self.pc[r][self.pos[r]] for r in live}
                self.errors.append(f"deadlock: {where}")
                break
            if self.errors:
                break
        # agreement of the two copies of every active bit at the end
        for k in range(self.K):
            for r in self.has[k]:
                if self.w_active[r][k] != self.s_active[k]:
                    self.errors.append(f"end: active bit of key {k} differs (worker {r})")
        return self.errors


def check(ranks=3, keys=12, blocks=4, passes=4, schedules=300, seed=0, mutate="", verbose=True) -> int:
    bad = 0
    for i in range(schedules):
        rng = random.Random(seed * 100003 + i)
        m = Model(ranks, keys, blocks, passes, revive=bool(i & 1), rng=rng, mutate=mutate)
        errs = m.run()
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
    ap.add_argument("--mutate", default="", choices=["", "no_reset", "no_republish"],
                    help="break the protocol on purpose: the checker must notice")
    a = ap.parse_args()
    n = check(a.ranks, a.keys, a.blocks, a.passes, a.schedules, mutate=a.mutate)
    print("violations:", n)
    raise SystemExit(1 if n else 0)
