"""Randomised interleaving check of the streaming runner's ordering (csrc/cuda/sgd_stream.cu):
minibatch k is uploaded and localised on copy lane k % L into ring set k % R with scratch k % L, and
stepped on the step stream; the only ordering between the streams are the events the runner
records and waits for.

It is a model of the enqueue order, not of the CUDA code: every stream is a FIFO of operations, an
operation may run when it is at the head of its stream and the events it waits for have fired, and
a seeded random scheduler picks among the runnable heads.  Checked on every schedule:
  * a ring set is never written (upload, localisation outputs) while a step still reads it, and a
    step only reads a set that holds ITS minibatch, completely localised;
  * a scratch set is never written by stage A of one minibatch while stage B of another still reads it;
  * the sketch (stage B) is updated in minibatch order;
  * no deadlock.

    python tools/runner_check.py [--ring 5] [--lanes 3] [--minibatches 40] [--schedules 300] [--mutate no_done]
"""
from __future__ import annotations

import argparse
import random


def program(n_mb: int, ring: int, mutate: str, lanes: int = 2):
    """The operations the runner enqueues, per stream, in order.  op = (kind, k, waits, records)."""
    streams = {f"copy{l}": [] for l in range(lanes)}
    streams["step"] = []
    used = [False] * ring
    prev_slot = None
    for k in range(n_mb):
        slot, lane = k % ring, k % lanes
        if mutate == "one_scratch":
            lane_scratch = 0
        else:
            lane_scratch = lane
        cs = streams[f"copy{lane}"]
        waits = []
        if used[slot] and mutate != "no_done":
            waits.append(("done", slot, k - ring))           # the step that last read this set
        cs.append(("upload", k, waits, [], slot, lane_scratch))
        cs.append(("stageA", k, [], [], slot, lane_scratch))
        waits_b = []
        if prev_slot is not None and mutate != "no_order":
            waits_b.append(("ready", prev_slot, k - 1))      # stage B of the previous minibatch
        cs.append(("stageB", k, waits_b, [("ready", slot, k)], slot, lane_scratch))
        streams["step"].append(("step", k, [("ready", slot, k)], [("done", slot, k)], slot, lane_scratch))
        used[slot] = True
        prev_slot = slot
    return streams


class Model:
    def __init__(self, n_mb, ring, rng, mutate="", lanes=2):
        self.rng, self.ring = rng, ring
        self.streams = program(n_mb, ring, mutate, lanes)
        self.pos = {s: 0 for s in self.streams}
        self.fired = set()                 # (event kind, slot, minibatch) as recorded
        self.set_holds = [None] * ring     # minibatch whose data the ring set holds (None: being written / empty)
        self.set_state = ["empty"] * ring  # empty | uploaded | sorted | ready | reading
        self.scratch_holds = [None] * lanes  # minibatch whose uniq / cnt the scratch set holds
        self.sketch_next = 0
        self.errors = []
        # operations are not atomic: model each as begin / end so that overlaps are visible
        self.running = {s: None for s in self.streams}

    def head(self, s):
        return self.streams[s][self.pos[s]] if self.pos[s] < len(self.streams[s]) else None

    def can_begin(self, s):
        op = self.head(s)
        return op is not None and self.running[s] is None and all(w in self.fired for w in op[2])

    def begin(self, s):
        kind, k, _, _, slot, scr = self.head(s)
        self.running[s] = (kind, k, slot, scr)
        if kind == "upload":
            if self.set_state[slot] == "reading":
                self.errors.append(f"minibatch {k}: upload into set {slot} while step {self.set_holds[slot]} reads it")
            self.set_holds[slot], self.set_state[slot] = k, "writing"
        elif kind == "stageA":
            for other in self.running.values():
                if other and other[0] == "stageB" and other[3] == scr and other[1] != k:
                    self.errors.append(f"minibatch {k}: stage A overwrites scratch {scr} under stage B of {other[1]}")
            self.scratch_holds[scr] = k
        elif kind == "stageB":
            if self.scratch_holds[scr] != k:
                self.errors.append(f"minibatch {k}: stage B reads scratch {scr} holding {self.scratch_holds[scr]}")
            if self.sketch_next != k:
                self.errors.append(f"minibatch {k}: sketch updated out of order (expected {self.sketch_next})")
            self.sketch_next = k + 1
        elif kind == "step":
            if self.set_holds[slot] != k or self.set_state[slot] != "ready":
                self.errors.append(f"step {k}: set {slot} holds {self.set_holds[slot]} in state {self.set_state[slot]}")
            self.set_state[slot] = "reading"

    def end(self, s):
        kind, k, slot, scr = self.running[s]
        op = self.head(s)
        if kind == "stageB":
            if self.scratch_holds[scr] != k:
                self.errors.append(f"minibatch {k}: scratch {scr} changed under stage B (now {self.scratch_holds[scr]})")
            if self.set_holds[slot] == k:
                self.set_state[slot] = "ready"
        elif kind == "step":
            if self.set_holds[slot] != k:
                self.errors.append(f"step {k}: set {slot} was overwritten while it ran (now {self.set_holds[slot]})")
            else:
                self.set_state[slot] = "idle"
        for ev in op[3]:
            self.fired.add(ev)
        self.running[s] = None
        self.pos[s] += 1

    def run(self):
        while True:
            moves = [("end", s) for s in self.streams if self.running[s] is not None] + \
                    [("begin", s) for s in self.streams if self.can_begin(s)]
            if not moves:
                if any(self.head(s) is not None for s in self.streams):
                    self.errors.append("deadlock: " + str({s: self.head(s)[:2] for s in self.streams if self.head(s)}))
                break
            what, s = self.rng.choice(moves)
            (self.end if what == "end" else self.begin)(s)
            if self.errors:
                break
        return self.errors


def check(ring=5, minibatches=40, schedules=300, seed=0, mutate="", verbose=True, lanes=2) -> int:
    bad = 0
    for i in range(schedules):
        errs = Model(minibatches, ring, random.Random(seed * 7919 + i), mutate, lanes).run()
        if errs:
            bad += 1
            if verbose:
                print(f"schedule {i}: {errs[0]}")
    return bad


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ring", type=int, default=5)
    ap.add_argument("--lanes", type=int, default=3)
    ap.add_argument("--minibatches", type=int, default=40)
    ap.add_argument("--schedules", type=int, default=300)
    ap.add_argument("--mutate", default="", choices=["", "no_done", "no_order", "one_scratch"],
                    help="break the ordering on purpose: the checker must notice")
    a = ap.parse_args()
    n = check(a.ring, a.minibatches, a.schedules, mutate=a.mutate, lanes=a.lanes)
    print("violations:", n)
    raise SystemExit(1 if n else 0)
