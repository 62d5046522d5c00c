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
