"""Executable model of the cross-GPU signalling protocol of one training step (CPU only).

compute-sanitizer's racecheck is closed on the GPU pool, so the LOGIC of the protocol
— which barriers exist, which were removed, why the one-shot staging area is double buffered —
is checked here on a small model: every rank is a sequence of atomic operations on versioned
memory cells, schedules interleave the ranks at random and adversarially (one rank runs as far
ahead as the protocol lets it), and every read asserts that it sees the version it is entitled
to.  Memory-model questions (fences, visibility) are outside this model; it assumes that a store
issued before a flag release is visible after the matching acquire, which is what the kernels'
`__threadfence_system(); st.release.sys` / `ld.acquire.sys` pairs provide.

Modelled step of rank r (csrc/comm/p2p_collectives.cu, csrc/reducer/reducer.cpp):
  backward, bucket by bucket: reads the bucket's weights (dgrad), then produces its gradients
  bucket b (two-shot, closing barrier skipped): opening barrier; read every peer's gradients of
      slice r; publish slice r of the new weights into every rank
  tail (one-shot push): push own gradients into slot [parity][r] of every peer's staging area;
      ONE barrier; reduce own staging slots; update own copy of the tail weights
  next step: clear own gradients (memset), forward reads ALL weights
The variants with a known hazard (single-buffered staging without any other collective in the
step; clearing gradients with no closing barrier anywhere) must be CAUGHT by the same model.
"""
import random

import pytest


class Violation(Exception):
    pass


class Rank:
    """Program of one rank as a list of (kind, payload) atomic operations."""

    def __init__(self, r, world, steps, buckets, double_buffer, tail, closing_barrier_on_last):
        self.r, self.world = r, world
        self.pc = 0
        self.epoch = 0            # barriers passed so far (device-side counter)
        self.uses = 0             # one-shot uses (double-buffer parity)
        ops = []
        for t in range(steps):
            ops.append(("clear_grads", t))
            ops.append(("forward", t))
            nb = buckets
            for b in range(nb):
                last = (b == nb - 1) and not tail
                ops.append(("backward", (t, b)))      # dgrad reads bucket b's weights, wgrad fills its gradients
                ops.append(("arrive", None))
                ops.append(("wait", None))
                ops.append(("pull", (t, b)))
                ops.append(("publish", (t, b)))
                if last and closing_barrier_on_last:
                    ops.append(("arrive", None))
                    ops.append(("wait", None))
            if tail:
                ops.append(("backward", (t, buckets)))
                ops.append(("push", (t, double_buffer)))
                ops.append(("arrive", None))
                ops.append(("wait", None))
                ops.append(("reduce", (t, double_buffer)))
        ops.append(("clear_grads", steps))      # the step after the last one starts normally
        ops.append(("forward", steps))
        self.ops = ops

    def done(self):
        return self.pc >= len(self.ops)


class World:
    def __init__(self, world, steps=3, buckets=2, double_buffer=True, tail=True,
                 closing_barrier_on_last=True):
        self.n = world
        self.buckets = buckets
        self.tail = tail
        self.ranks = [Rank(r, world, steps, buckets, double_buffer, tail, closing_barrier_on_last)
                      for r in range(world)]
        # versioned memory: value = the step whose data the cell holds (None = cleared / nothing)
        self.grads = [[None] * (buckets + 1) for _ in range(world)]           # [rank][bucket|tail]
        self.param = [[[0] * world for _ in range(buckets)] for _ in range(world)]   # [rank][bucket][slice]
        self.tail_param = [0] * world
        self.staging = [[[None] * world for _ in range(2)] for _ in range(world)]   # [rank][parity][src]
        self.flags = [[0] * world for _ in range(world)]                       # flags[dst][src]

    def runnable(self, rk):
        if rk.done():
            return False
        kind, _ = rk.ops[rk.pc]
        if kind == "wait":
            return all(self.flags[rk.r][p] >= rk.epoch for p in range(self.n))
        return True

    def step(self, rk):
        kind, arg = rk.ops[rk.pc]
        r, n = rk.r, self.n
        if kind == "clear_grads":
            self.grads[r] = [None] * (self.buckets + 1)
        elif kind == "forward":
            t = arg
            for b in range(self.buckets):
                for s in range(n):
                    if self.param[r][b][s] != t:
                        raise Violation(f"rank {r} forward of step {t}: bucket {b} slice {s} is at "
                                        f"version {self.param[r][b][s]}")
            if self.tail and self.tail_param[r] != t:
                raise Violation(f"rank {r} forward of step {t}: tail weights at {self.tail_param[r]}")
        elif kind == "backward":
            t, b = arg
            if b < self.buckets:
                for s_ in range(n):        # a peer must not have published step t+1 weights yet
                    if self.param[r][b][s_] != t:
                        raise Violation(f"rank {r} backward of step {t}: bucket {b} slice {s_} weights "
                                        f"already at {self.param[r][b][s_]}")
            elif self.tail_param[r] != t:
                raise Violation(f"rank {r} backward of step {t}: tail weights at {self.tail_param[r]}")
            self.grads[r][b] = t
        elif kind == "arrive":
            rk.epoch += 1
            for p in range(n):
                self.flags[p][r] = rk.epoch
        elif kind == "wait":
            pass                                   # runnable() already checked the condition
        elif kind == "pull":
            t, b = arg
            for p in range(n):
                if self.grads[p][b] != t:
                    raise Violation(f"rank {r} step {t} bucket {b}: rank {p}'s gradients hold {self.grads[p][b]}")
        elif kind == "publish":
            t, b = arg
            for p in range(n):
                self.param[p][b][r] = t + 1
        elif kind == "push":
            t, dbl = arg
            par = rk.uses & 1 if dbl else 0
            if self.grads[r][self.buckets] != t:
                raise Violation(f"rank {r} step {t}: own tail gradients hold {self.grads[r][self.buckets]}")
            for p in range(n):
                self.staging[p][par][r] = t
        elif kind == "reduce":
            t, dbl = arg
            par = rk.uses & 1 if dbl else 0
            for p in range(n):
                if self.staging[r][par][p] != t:
                    raise Violation(f"rank {r} step {t}: staging slot of rank {p} holds {self.staging[r][par][p]}")
            self.tail_param[r] = t + 1
            rk.uses += 1
        rk.pc += 1

    def run(self, pick):
        while True:
            ready = [rk for rk in self.ranks if self.runnable(rk)]
            if not ready:
                if all(rk.done() for rk in self.ranks):
                    return
                raise Violation("deadlock: " + ", ".join(str(rk.ops[rk.pc]) for rk in self.ranks if not rk.done()))
            self.step(pick(ready))


def schedules(seed, n_random):
    rng = random.Random(seed)
    for _ in range(n_random):
        yield lambda ready: rng.choice(ready)
    # adversarial: a fixed priority order — the first rank always runs as far as it can
    for order in range(6):
        yield lambda ready, o=order: sorted(ready, key=lambda rk: (rk.r * (o + 1)) % 7)[0]
        yield lambda ready, o=order: sorted(ready, key=lambda rk: -((rk.r * (o + 1)) % 7))[0]


def explore(n_random=400, **kw):
    for world in (2, 3, 4):
        for pick in schedules(world, n_random):
            World(world, **kw).run(pick)


def test_step_protocol_as_built_has_no_hazard():
    """Two-shot buckets without closing barriers + double-buffered one-shot tail: every pull sees
    the peers' gradients of THIS step, every forward sees fully updated weights, every reduce
    sees this step's staging data — under random and adversarial schedules, 2-4 ranks."""
    explore(steps=3, buckets=2, double_buffer=True, tail=True)
    explore(steps=3, buckets=1, double_buffer=True, tail=True)


def test_last_bucket_keeps_its_closing_barrier_when_there_is_no_tail():
    explore(steps=3, buckets=2, tail=False, closing_barrier_on_last=True)


def test_model_catches_a_step_with_no_closing_barrier_at_all():
    """If NO collective of a step kept a closing barrier, a fast rank would clear its gradients
    (or start its forward) while a slow peer still needs them: the model must find that."""
    with pytest.raises(Violation):
        explore(steps=3, buckets=2, tail=False, closing_barrier_on_last=False)


def test_single_buffered_staging_is_safe_only_behind_another_collective():
    """Why the staging area is double buffered: with one buffer and the tail as the step's ONLY
    collective a fast rank pushes step t+1 into slots a slow rank has not reduced yet; with other
    bucket collectives in between the single buffer would do — the kernel does not rely on that."""
    with pytest.raises(Violation):
        explore(steps=3, buckets=0, double_buffer=False, tail=True)
    explore(steps=3, buckets=0, double_buffer=True, tail=True)
    explore(steps=3, buckets=1, double_buffer=False, tail=True)
