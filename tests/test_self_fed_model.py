"""Randomised model of the SELF-FED pipeline of the two-step kernel (``csrc/temporal2d.cu``, "self-fed pipeline"):
no producer warp, no "empty" barriers -- thread 0 requests T block q+3 and Cp block q+2 right after the per-block
barrier of block q, and the Cp row that step 2 of block q needs from block q-1 is carried in registers.

Actors: W consumer warps (warp 0 also issues the loads) and the TMA engine (loads complete after a random delay, in
any order).  Checked on every read: the slot holds the block the reader expects (no overwrite before the last read, no
read before arrival); the run must terminate (no deadlock).  Ring depths are read from the kernel's header.  A
schedule that requests too early (e.g. Cp q+2 without the carried row, or T q+4) must be caught.
"""
import os
import random
import re

import pytest

HDR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "csrc", "temporal2d_core.h")


def ring_depths():
    src = open(HDR).read()
    return tuple(int(re.search(rf"constexpr int {n} = (\d+);", src).group(1)) for n in ("SNT", "SNC", "SNM"))


def simulate(nblocks, nwarps, seed, t_ahead=3, c_ahead=2, carry_cp_row=True):
    SNT, SNC, SNM = ring_depths()
    rng = random.Random(seed)
    slotT, slotC, slotM = [None] * SNT, [None] * SNC, [None] * SNM     # which block each slot holds (None = in flight / empty)
    inflight = []                                                      # (ring, slot, block)
    issuedT = issuedC = 0

    def issue(ring):
        nonlocal issuedT, issuedC
        if ring == "T" and issuedT < nblocks:
            slotT[issuedT % SNT] = None
            inflight.append(("T", issuedT % SNT, issuedT))
            issuedT += 1
        if ring == "C" and issuedC < nblocks:
            slotC[issuedC % SNC] = None
            inflight.append(("C", issuedC % SNC, issuedC))
            issuedC += 1

    for _ in range(t_ahead):
        issue("T")
    for _ in range(c_ahead):
        issue("C")
    # warp program: per block q: wait T(q), T(q+1), C(q); step 1 (reads T q-1, q, q+1, C q; writes M q); barrier;
    # [warp 0: issue]; step 2 (reads M q-1, q; C q and, without the carried row, C q-1)
    pc = [(0, "wait")] * nwarps
    at_barrier = set()
    steps = 0
    while any(q < nblocks for q, _ in pc):
        steps += 1
        assert steps < 200000, "no progress: deadlock"
        if inflight and rng.random() < 0.4:                            # a load lands
            ring, slot, blk = inflight.pop(rng.randrange(len(inflight)))
            (slotT if ring == "T" else slotC)[slot] = blk
            continue
        w = rng.randrange(nwarps)
        q, st = pc[w]
        if q >= nblocks:
            continue
        if st == "wait":
            need = [("T", q), ("C", q)] + ([("T", q + 1)] if q + 1 < nblocks else [])
            if all((slotT if r == "T" else slotC)[b % (SNT if r == "T" else SNC)] == b for r, b in need):
                pc[w] = (q, "step1")
        elif st == "step1":
            for b in ([q - 1] if q > 0 else []) + [q] + ([q + 1] if q + 1 < nblocks else []):
                assert slotT[b % SNT] == b, f"step 1 of block {q}: T slot holds {slotT[b % SNT]}, wanted {b}"
            assert slotC[q % SNC] == q
            slotM[q % SNM] = q
            pc[w] = (q, "barrier")
            at_barrier.add(w)
        elif st == "barrier":
            if len(at_barrier) == nwarps:
                for v in range(nwarps):
                    pc[v] = (pc[v][0], "issue" if v == 0 else "step2")
                at_barrier.clear()
        elif st == "issue":
            issue("T")
            issue("C")
            pc[w] = (q, "step2")
        elif st == "step2":
            for b in ([q - 1] if q > 0 else []) + [q]:
                assert slotM[b % SNM] == b, f"step 2 of block {q}: mid slot holds {slotM[b % SNM]}, wanted {b}"
            assert slotC[q % SNC] == q, f"step 2 of block {q}: Cp slot holds {slotC[q % SNC]}"
            if not carry_cp_row and q > 0:
                assert slotC[(q - 1) % SNC] == q - 1, f"step 2 of block {q}: Cp block {q - 1} already overwritten"
            pc[w] = (q + 1, "wait")
    return True


@pytest.mark.parametrize("nwarps", [2, 4, 8])
def test_self_fed_schedule_is_safe_and_live(nwarps):
    for seed in range(25):
        assert simulate(nblocks=40, nwarps=nwarps, seed=seed)


def test_model_catches_a_missing_carried_cp_row():
    with pytest.raises(AssertionError):
        for seed in range(50):
            simulate(nblocks=40, nwarps=4, seed=seed, carry_cp_row=False)


def test_model_catches_requests_that_run_too_far_ahead():
    with pytest.raises(AssertionError):
        for seed in range(50):
            simulate(nblocks=40, nwarps=4, seed=seed, t_ahead=4)


# ---- mid lines of a y-halo column: when may a consumer read them? ------------------------------------------------------
def simulate_midline_reads(seed, read_after_first_block):
    """One CTA starting an item that lies on a y face.  Actors: the NEIGHBOUR (writes the mid line, then releases its
    flag), THREAD 0 (acquires the flag, then requests the item's first T block), the TMA engine (the block arrives some
    time after the request) and a CONSUMER thread that prefetches the mid-line values of the item's first rows --
    either as soon as it reaches the item (the bug found on 4 GPUs) or only after the first T block has arrived (the
    fix: the arrival is ordered behind thread 0's acquire).  Returns True if the consumer saw the new line."""
    rng = random.Random(seed)
    line_new = flag = requested = arrived = False
    consumer_value = None
    consumer_at_item = False
    for _ in range(10000):
        actor = rng.choice(["neighbour", "thread0", "tma", "consumer"])
        if actor == "neighbour":
            if not line_new:
                line_new = True                      # stores of the mid line (then fence + release)
            elif not flag:
                flag = True
        elif actor == "thread0":
            if flag and not requested:               # ld.acquire saw the flag: request the first T block
                requested = True
        elif actor == "tma":
            if requested and not arrived and rng.random() < 0.5:
                arrived = True
        elif actor == "consumer" and consumer_value is None:
            consumer_at_item = True
            if read_after_first_block and not arrived:
                continue                             # mbarrier wait on the first T block
            consumer_value = "new" if line_new else "stale"
        if consumer_value is not None:
            break
    assert consumer_at_item
    return consumer_value == "new"


def test_midline_prefetch_is_safe_only_behind_the_first_block():
    assert all(simulate_midline_reads(seed, read_after_first_block=True) for seed in range(300))
    assert not all(simulate_midline_reads(seed, read_after_first_block=False) for seed in range(300)), \
        "the model must reproduce the early read that bench.py's verify block caught"
