"""Randomised model of the producer / consumer pipeline inside the two-steps-per-pass kernel.

The host emulator (``tests/test_temporal_core.py``) runs the kernel's index code but executes the producer's
loads just in time and the consumers one after the other, so it cannot see whether the ring depths and the
places where slots are handed back are right.  This model does: the TMA producer, the asynchronous TMA
completions and every consumer warp are separate actors under a random scheduler, mbarriers have real phase
parity semantics, and each slot carries a label saying which block it holds.  Checked:

* a consumer only ever reads a slot that holds exactly the block it expects (never an older or newer one,
  never one with a TMA write in flight);
* the producer only starts a TMA into a slot after every warp has finished reading its previous content;
* a warp only overwrites its part of a mid-ring slot after every warp has finished the step-2 reads of the
  previous content (the single consumer barrier per block has to guarantee that);
* no deadlock.

The actor programs mirror ``diff2d_temporal2_kernel`` in ``csrc/temporal2d.cu`` statement by statement (ring
depths are parameters so that the test can also show that shallower rings break).  The reference has nothing
comparable (no native code; SURVEY.md section 5: no race detection).
"""
import random

import pytest


class MBar:
    """mbarrier with `count` arrivals per phase; `done` = number of completed phases."""

    def __init__(self, count):
        self.count, self.pending, self.done = count, count, 0

    def arrive(self):
        self.pending -= 1
        if self.pending == 0:
            self.pending = self.count
            self.done += 1

    def passes(self, parity):
        # try_wait.parity(p): true once the phase with parity p has completed, i.e. the current phase has the
        # other parity (a fresh barrier passes parity 1)
        return (self.done & 1) != parity


class Model:
    def __init__(self, items, ncw, snt, snc, snm, seed):
        self.items, self.ncw, self.snt, self.snc, self.snm = items, ncw, snt, snc, snm
        self.rng = random.Random(seed)
        self.fullT = [MBar(1) for _ in range(snt)]
        self.emptyT = [MBar(ncw) for _ in range(snt)]
        self.fullC = [MBar(1) for _ in range(snc)]
        self.emptyC = [MBar(ncw) for _ in range(snc)]
        self.slotT = [None] * snt          # label (global block counter) or "flux"
        self.slotC = [None] * snc
        self.slotM = [[None] * ncw for _ in range(snm)]   # per-warp parts
        self.readsT = [[-1] * ncw for _ in range(snt)]    # last label each warp finished reading from the slot
        self.readsC = [[-1] * ncw for _ in range(snc)]
        self.step2_done = [-1] * ncw                      # global block counter of the last finished step 2
        self.bar_gen, self.bar_arrived = 0, 0
        self.inflight = []                                # pending TMA completions: (ring, slot, label)

    # ---- actors (generators yield predicates to wait for; `None` = just a scheduling point) ----------------
    def producer(self):
        tq = cq = 0
        for nbi in self.items:
            def issue(ring, q, full, empty, slots, reads, depth):
                s = q % depth
                if q >= depth:
                    par = ((q // depth) - 1) & 1
                    yield lambda: empty[s].passes(par)
                    # every warp must be done with the previous content (label q - depth)
                    assert all(r >= q - depth for r in reads[s]), \
                        f"producer refills {ring} slot {s} (block {q}) while block {q - depth} is still being read"
                assert slots[s] != "flux"
                slots[s] = "flux"
                self.inflight.append((ring, s, q))
                yield None
            yield from issue("T", tq, self.fullT, self.emptyT, self.slotT, self.readsT, self.snt)
            tq += 1
            for b in range(nbi):
                if b + 1 < nbi:
                    yield from issue("T", tq, self.fullT, self.emptyT, self.slotT, self.readsT, self.snt)
                    tq += 1
                yield from issue("C", cq, self.fullC, self.emptyC, self.slotC, self.readsC, self.snc)
                cq += 1

    def tma(self):
        """Completes in-flight loads in any order."""
        while True:
            yield lambda: bool(self.inflight)
            ring, s, q = self.inflight.pop(self.rng.randrange(len(self.inflight)))
            if ring == "T":
                self.slotT[s] = q
                self.fullT[s].arrive()
            else:
                self.slotC[s] = q
                self.fullC[s].arrive()

    def consumer(self, w):
        snt, snc, snm = self.snt, self.snc, self.snm
        tq = cq = mq = 0

        def readT(q):
            assert self.slotT[q % snt] == q, f"warp {w} expected T block {q}, slot holds {self.slotT[q % snt]}"

        def readC(q):
            assert self.slotC[q % snc] == q, f"warp {w} expected Cp block {q}, slot holds {self.slotC[q % snc]}"

        for nbi in self.items:
            yield lambda: self.fullT[tq % snt].passes((tq // snt) & 1)
            for b in range(nbi):
                qb, qc, qm = tq + b, cq + b, mq + b
                if b + 1 < nbi:
                    yield lambda: self.fullT[(qb + 1) % snt].passes(((qb + 1) // snt) & 1)
                yield lambda: self.fullC[qc % snc].passes((qc // snc) & 1)
                # ---- step 1: reads T blocks b-1, b, b+1 and Cp block b, writes its part of mid block b
                if b > 0:
                    readT(qb - 1)
                readT(qb)
                if b + 1 < nbi:
                    readT(qb + 1)
                readC(qc)
                yield None
                # the slot's previous content, block qm - snm, is read by step 2 of the blocks qm-snm and qm-snm+1
                if qm >= snm:
                    assert all(d >= qm - snm + 1 for d in self.step2_done), \
                        f"warp {w} overwrites mid block {qm - snm} before every warp finished reading it"
                self.slotM[qm % snm][w] = qm
                if b > 0:
                    self.readsT[(qb - 1) % snt][w] = qb - 1
                    self.emptyT[(qb - 1) % snt].arrive()
                # ---- consumer barrier
                gen = self.bar_gen
                self.bar_arrived += 1
                if self.bar_arrived == self.ncw:
                    self.bar_arrived = 0
                    self.bar_gen += 1
                yield lambda: self.bar_gen > gen
                # ---- step 2: reads mid blocks b-1 (b > 0) and b from all warps, Cp blocks b-1 and b
                assert all(p == qm for p in self.slotM[qm % snm]), f"warp {w}: mid block {qm} incomplete"
                if b > 0:
                    assert all(p == qm - 1 for p in self.slotM[(qm - 1) % snm]), f"warp {w}: mid block {qm - 1} gone"
                    readC(qc - 1)
                readC(qc)
                yield None
                self.step2_done[w] = qm
                if b > 0:
                    self.readsC[(qc - 1) % snc][w] = qc - 1
                    self.emptyC[(qc - 1) % snc].arrive()
            self.readsT[(tq + nbi - 1) % snt][w] = tq + nbi - 1
            self.emptyT[(tq + nbi - 1) % snt].arrive()
            self.readsC[(cq + nbi - 1) % snc][w] = cq + nbi - 1
            self.emptyC[(cq + nbi - 1) % snc].arrive()
            tq += nbi
            cq += nbi
            mq += nbi

    # ---- scheduler -----------------------------------------------------------------------------------------
    def run(self):
        actors = {"P": self.producer(), "TMA": self.tma()}
        for w in range(self.ncw):
            actors[f"W{w}"] = self.consumer(w)
        waiting = {}
        for name, gen in list(actors.items()):
            try:
                waiting[name] = next(gen)
            except StopIteration:
                del actors[name]
        while any(n != "TMA" for n in actors):
            ready = [n for n in actors if waiting[n] is None or waiting[n]()]
            assert ready, f"deadlock; still alive: {sorted(actors)}"
            n = self.rng.choice(ready)
            try:
                waiting[n] = next(actors[n])
            except StopIteration:
                del actors[n]
                del waiting[n]


def _ring_depths():
    """The depths the kernel is actually built with (csrc/temporal2d_core.h)."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "csrc",
                            "temporal2d_core.h")).read()
    return tuple(int(re.search(rf"constexpr int {n} = (\d+);", src).group(1)) for n in ("SNT", "SNC", "SNM"))


def _items(rng, n):
    return [rng.choice([1, 1, 2, 3, 5, 8, 16]) for _ in range(n)]


@pytest.mark.parametrize("ncw", [2, 4, 16])
def test_pipeline_of_the_two_step_kernel(ncw):
    rng = random.Random(ncw)
    snt, snc, snm = _ring_depths()
    assert (snt, snc, snm) == (4, 3, 3)
    for seed in range(30):
        Model(_items(rng, 12), ncw, snt=snt, snc=snc, snm=snm, seed=seed).run()


@pytest.mark.parametrize("depths", [(3, 3, 3), (4, 2, 3)])
def test_one_slot_less_still_works_but_cannot_prefetch(depths):
    """With one slot less in the T or Cp ring the hand-back points still make the pipeline correct (the producer
    just cannot run ahead any more): the depths 4 / 3 are a performance choice, not a correctness requirement."""
    snt, snc, snm = depths
    rng = random.Random(3)
    for seed in range(20):
        Model(_items(rng, 10), 4, snt=snt, snc=snc, snm=snm, seed=seed).run()


@pytest.mark.parametrize("depths", [(2, 3, 3), (4, 1, 3), (4, 3, 2)])
def test_too_shallow_rings_are_caught(depths):
    """T ring of 2: block b+1 cannot be loaded before b-1 is handed back, which needs b+1 -> deadlock.  Cp ring
    of 1: same with blocks b and b-1.  Mid ring of 2: a warp one barrier ahead overwrites the block a slower
    warp still reads as `prev`."""
    snt, snc, snm = depths
    rng = random.Random(7)
    with pytest.raises(AssertionError):
        for seed in range(200):
            Model(_items(rng, 12), 4, snt=snt, snc=snc, snm=snm, seed=seed).run()
