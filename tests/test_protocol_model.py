"""Randomised model of the cross-GPU epoch-flag protocol (no GPU, no processes).

``compute-sanitizer --tool racecheck`` cannot see traffic between GPUs, and the reference has no race
detection at all.  This test executes the protocol of ``csrc/push.cuh`` /
``csrc/stencil2d.cu`` as an abstract machine under a random scheduler and checks the two properties the
kernels rely on:

* RAW: when an edge tile of step ``k`` reads a halo plane, that plane holds exactly the neighbour's step
  ``k-1`` values (not older, not newer);
* WAR: when a rank pushes step ``k`` values into a neighbour's buffer, no tile of the neighbour still has to
  read the previous contents of those halo cells.

Model: ranks on a periodic ring (one decomposed dimension), two field buffers per rank that alternate,
per rank a "low" and a "high" edge made of several tiles plus interior tiles.  A tile of step ``k``
(a) if it is an edge tile, waits until the facing neighbour's flag is >= ``k-1`` and then reads the halo
version, (b) computes, (c) pushes its edge cells into the neighbour's *new* buffer, (d) bumps the rank's
private arrival counter; the last tile of an edge releases flag ``k``.  Kernels of one rank run one after
the other (stream order) but tiles inside a kernel run in any order, and ranks drift freely.
"""
import random

import pytest


class Rank:
    def __init__(self, r, ntiles_edge, ntiles_interior):
        self.r = r
        self.step = 1                          # kernel currently executing (steps start at 1)
        # halo version stored in buffer b on side s: halo[b][s]; step k reads buffer (k-1)%2 ... see below
        self.halo = [[0, 0], [0, 0]]           # begin(): both buffers consistent at "epoch 0"
        self.flag = [0, 0]                     # flag[s]: last epoch released BY the neighbour on side s
        self.count = [0, 0]                    # private arrival counters per outgoing side
        self.nte, self.nti = ntiles_edge, ntiles_interior
        self.todo = self._tiles()

    def _tiles(self):
        # (kind, side, phase): edge tiles go through phases wait -> push; interior tiles are one phase
        t = [("edge", s, "wait") for s in (0, 1) for _ in range(self.nte)]
        t += [("interior", None, "run")] * self.nti
        return t


def simulate(nranks, nsteps, nte, nti, seed, frame_first):
    rng = random.Random(seed)
    ranks = [Rank(r, nte, nti) for r in range(nranks)]
    ops = 0
    while any(rk.step <= nsteps for rk in ranks):
        # pick a rank that still has work and an enabled tile operation
        cand = []
        for rk in ranks:
            if rk.step > nsteps:
                continue
            k = rk.step
            idxs = range(len(rk.todo))
            if frame_first:                    # persistent kernel: tiles are taken in list order, a few in flight
                idxs = range(min(len(rk.todo), 3))
            for i in idxs:
                kind, side, phase = rk.todo[i]
                if kind == "edge" and phase == "wait" and rk.flag[side] < k - 1:
                    continue                   # blocked on the neighbour's previous epoch
                cand.append((rk, i))
        assert cand, "deadlock: no enabled operation"
        rk, i = rng.choice(cand)
        k = rk.step
        kind, side, phase = rk.todo[i]
        old_buf, new_buf = (k - 1) % 2, k % 2  # step k reads buffer (k-1)%2 and writes buffer k%2
        if kind == "interior":
            rk.todo.pop(i)
        elif phase == "wait":
            # RAW check: the halo this tile reads must be exactly the neighbour's step k-1 push
            assert rk.halo[old_buf][side] == k - 1, (
                f"RAW violation: rank {rk.r} step {k} side {side} read halo version {rk.halo[old_buf][side]}")
            rk.todo[i] = (kind, side, "push")
        else:
            nb = ranks[(rk.r + (1 if side == 1 else -1)) % nranks]
            nb_side = 1 - side                 # my high edge lands in the neighbour's low halo and vice versa
            # WAR check: the neighbour must not still need the old contents of buffer new_buf on that side,
            # i.e. it must have finished (passed the wait phase of) every edge tile of the step that reads it.
            # Buffer new_buf was last read by the neighbour's step k-1 (it reads buffer (k-2)%2 == k%2).
            if nb.step < k - 1:
                raise AssertionError(f"WAR violation: rank {rk.r} step {k} overwrites halo that neighbour "
                                     f"{nb.r} (at step {nb.step}) has not read yet")
            if nb.step == k - 1:
                pending = [t for t in nb.todo if t[0] == "edge" and t[1] == nb_side and t[2] == "wait"]
                assert not pending, (f"WAR violation: rank {rk.r} step {k} overwrites halo still to be read by "
                                     f"{len(pending)} tile(s) of rank {nb.r} step {nb.step}")
            # the halo cell keeps the newest version; all tiles of an edge write disjoint cells, so the
            # plane is complete (version k) once the flag says so -- modelled at release time below
            rk.count[side] += 1
            if rk.count[side] % rk.nte == 0:   # last tile of this edge: release the epoch
                nb.halo[new_buf][nb_side] = k
                assert nb.flag[nb_side] == k - 1, "epochs must be released in order"
                nb.flag[nb_side] = k
            rk.todo.pop(i)
        ops += 1
        if not rk.todo:                        # kernel finished: next launch in stream order
            rk.step += 1
            rk.todo = rk._tiles()
    return ops


@pytest.mark.parametrize("nranks", [1, 2, 3, 5])
@pytest.mark.parametrize("frame_first", [True, False])
def test_epoch_protocol_is_race_and_deadlock_free(nranks, frame_first):
    for seed in range(40):
        simulate(nranks, nsteps=12, nte=3, nti=4, seed=seed, frame_first=frame_first)


def test_model_catches_a_broken_protocol():
    """Sanity of the checker itself: waiting for epoch k-2 instead of k-1 must be flagged."""
    def broken(nranks, seed):
        rng = random.Random(seed)
        ranks = [Rank(r, 2, 1) for r in range(nranks)]
        # same loop as simulate() but with the wait weakened to k-2
        for _ in range(20000):
            cand = []
            for rk in ranks:
                if rk.step > 8:
                    continue
                for i, (kind, side, phase) in enumerate(rk.todo):
                    if kind == "edge" and phase == "wait" and rk.flag[side] < rk.step - 2:
                        continue
                    cand.append((rk, i))
            if not cand:
                return
            rk, i = rng.choice(cand)
            k = rk.step
            kind, side, phase = rk.todo[i]
            if kind == "interior":
                rk.todo.pop(i)
            elif phase == "wait":
                assert rk.halo[(k - 1) % 2][side] == k - 1
                rk.todo[i] = (kind, side, "push")
            else:
                nb = ranks[(rk.r + (1 if side == 1 else -1)) % nranks]
                rk.count[side] += 1
                if rk.count[side] % rk.nte == 0:
                    nb.halo[k % 2][1 - side] = k
                    nb.flag[1 - side] = max(nb.flag[1 - side], k)
                rk.todo.pop(i)
            if not rk.todo:
                rk.step += 1
                rk.todo = rk._tiles()

    with pytest.raises(AssertionError):
        for seed in range(200):
            broken(3, seed)


# ======================================================================================================
# Two-steps-per-pass launches mixed with one-step launches (csrc/temporal2d.cu)
# ======================================================================================================
# A pass of epoch e is two kernels in stream order: F(e) ("mid lines": every CTA first waits for the DATA flags
# e-1 of BOTH neighbours, then stores its part of the rank's step-(k+1) edge line into the neighbour's mid
# buffer; the last CTA of a face releases the neighbour's mid flag e) and B(e) (edge items wait for the mid flag
# e of their side, read the mid line and the halo, push the step-(k+2) edge into the neighbour's new buffer;
# the last one releases DATA e).  B(e) itself never waits for DATA e-1: F(e) did, earlier on the same stream.
# A one-step launch S(e) behaves as in the model above.  One epoch per launch, buffers alternate per launch.
class Rank2:
    def __init__(self, r, sched, nte, nti, nf):
        self.r, self.nte, self.nti, self.nf = r, nte, nti, nf
        self.kernels = []
        for e, s in enumerate(sched, start=1):
            self.kernels += [("S", e)] if s == "S" else [("F", e), ("B", e)]
        self.k = 0
        self.halo = [[0, 0], [0, 0]]
        self.mid = [0, 0]
        self.data_flag = [0, 0]
        self.mid_flag = [0, 0]
        self.cnt = {"S": [0, 0], "B": [0, 0], "F": [0, 0]}     # the kernels keep separate arrival counters
        self.todo = self._tiles()

    def _tiles(self):
        if self.k >= len(self.kernels):
            return []
        kind, _ = self.kernels[self.k]
        if kind == "F":
            return [("face", s, "wait") for s in (0, 1) for _ in range(self.nf)]
        return [("edge", s, "wait") for s in (0, 1) for _ in range(self.nte)] + [("interior", None, "run")] * self.nti

    def done(self):
        return self.k >= len(self.kernels)

    def pending_readers(self, what, side, e):
        """Does this rank still have to read `what` ('mid' or the halo of buffer e % 2) on `side` in a launch
        older than e?"""
        for idx in range(self.k, len(self.kernels)):
            kind, ee = self.kernels[idx]
            if ee >= e:
                break
            todo = self.todo if idx == self.k else None          # None = not started: everything pending
            if what == "mid":
                if kind == "F":
                    return True                                  # its B(ee) comes later and reads the mid line
                if kind == "B" and (todo is None or any(t == ("edge", side, "wait") for t in todo)):
                    return True
            else:
                if (ee - 1) % 2 != e % 2:
                    continue                                     # that launch reads the other buffer
                if kind == "F":
                    if todo is None or any(t[0] == "face" and t[2] == "wait" for t in todo):
                        return True                              # F CTAs read the halos of both sides
                elif todo is None or any(t == ("edge", side, "wait") for t in todo):
                    return True
        return False


def simulate_mixed(nranks, sched, nte, nti, nf, seed, weaken=None):
    rng = random.Random(seed)
    ranks = [Rank2(r, sched, nte, nti, nf) for r in range(nranks)]
    while not all(rk.done() for rk in ranks):
        cand = []
        for rk in ranks:
            if rk.done():
                continue
            kind, e = rk.kernels[rk.k]
            for i, (tk, side, phase) in enumerate(rk.todo):
                if phase == "wait":
                    if kind == "S" and rk.data_flag[side] < e - 1:
                        continue
                    if kind == "B" and rk.mid_flag[side] < (e - 1 if weaken == "mid" else e):
                        continue
                    if kind == "F" and weaken != "data" and min(rk.data_flag) < e - 1:
                        continue
                cand.append((rk, i))
        assert cand, "deadlock: no enabled operation"
        rk, i = rng.choice(cand)
        kind, e = rk.kernels[rk.k]
        tk, side, phase = rk.todo[i]
        old_buf, new_buf = (e - 1) % 2, e % 2
        if tk == "interior":
            rk.todo.pop(i)
        elif phase == "wait":
            if kind == "F":
                assert rk.halo[old_buf] == [e - 1, e - 1], f"RAW: F({e}) on rank {rk.r} read halos {rk.halo[old_buf]}"
                rk.todo[i] = (tk, side, "store")
            else:
                assert rk.halo[old_buf][side] == e - 1, f"RAW: {kind}({e}) rank {rk.r} side {side} halo {rk.halo[old_buf][side]}"
                if kind == "B":
                    assert rk.mid[side] == e, f"RAW: B({e}) rank {rk.r} side {side} read mid version {rk.mid[side]}"
                rk.todo[i] = (tk, side, "push")
        else:
            nb = ranks[(rk.r + (1 if side == 1 else -1)) % nranks]
            nb_side = 1 - side
            if phase == "store":
                assert not nb.pending_readers("mid", nb_side, e), \
                    f"WAR: F({e}) of rank {rk.r} overwrites a mid line rank {nb.r} has not read yet"
                rk.cnt["F"][side] += 1
                if rk.cnt["F"][side] % rk.nf == 0:
                    assert nb.mid_flag[nb_side] < e
                    nb.mid[nb_side] = e
                    nb.mid_flag[nb_side] = e
            else:
                assert not nb.pending_readers("halo", nb_side, e), \
                    f"WAR: {kind}({e}) of rank {rk.r} overwrites a halo rank {nb.r} has not read yet"
                rk.cnt[kind][side] += 1
                if rk.cnt[kind][side] % rk.nte == 0:
                    assert nb.data_flag[nb_side] == e - 1, "epochs must be released in order"
                    nb.halo[new_buf][nb_side] = e
                    nb.data_flag[nb_side] = e
            rk.todo.pop(i)
        if not rk.todo:
            rk.k += 1
            rk.todo = rk._tiles()


@pytest.mark.parametrize("nranks", [1, 2, 3, 5])
@pytest.mark.parametrize("sched", ["PPPPPP", "PSPSPP", "SSPPSS", "PPPS", "SPSPSPS"])
def test_two_step_passes_are_race_and_deadlock_free(nranks, sched):
    for seed in range(25):
        simulate_mixed(nranks, sched, nte=3, nti=3, nf=2, seed=seed)


@pytest.mark.parametrize("weaken", ["mid", "data"])
def test_mixed_model_catches_broken_waits(weaken):
    """The checker must notice a B that accepts the previous pass's mid flag, and an F that does not wait for
    the neighbours' DATA flags."""
    with pytest.raises(AssertionError):
        for seed in range(300):
            simulate_mixed(3, "PPPPPP", nte=2, nti=1, nf=2, seed=seed, weaken=weaken)
