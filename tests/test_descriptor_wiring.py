"""Flag / buffer wiring of the push and mid-line descriptors, checked with a mock fabric (no GPU).

Every rank builds its descriptors independently; they only work if rank A's "slot I release in B's pad" is the
very slot B polls for A.  A mock fabric hands out fake addresses so that the wiring of
``PushPlan._build`` / ``PushPlan.mid_lines`` can be cross-checked between all ranks of a layout.
(The reference leaves all of this to MPI message matching inside ImplicitGlobalGrid.)
"""
import itertools

import pytest
import torch

from rocm_mpi_b200.ops import native
from rocm_mpi_b200.parallel.fabric import PushPlan
from rocm_mpi_b200.parallel.topology import PROC_NULL, CartTopology

PAD_BYTES = native.FLAG_BANKS * native.BANK_STRIDE * 8


class MockFabric:
    """Per-rank view: flag pad of rank r at 0x1000_0000 * (r+1), counter pad at +0x800_0000, fields elsewhere."""

    def __init__(self, me, dims, periods):
        self.me = me
        n = dims[0] * dims[1]
        self.topo = CartTopology.create(n, me, (dims[0], dims[1], 1), (periods[0], periods[1], 0))
        self.neighbors = []
        for o in itertools.product((-1, 0, 1), (-1, 0, 1), (0,)):
            if o != (0, 0, 0):
                r = self.topo.rank_at_offset(o)
                if r != PROC_NULL:
                    self.neighbors.append((o, r))
        self._next_field = 0
        self._fields = {}

    def bank_ptr(self, bank, rank=None):
        r = self.me if rank is None else rank
        return 0x10000000 * (r + 1) + bank * native.BANK_STRIDE * 8

    def count_bank_ptr(self, bank):
        return 0x10000000 * (self.me + 1) + 0x8000000 + bank * native.BANK_STRIDE * 8

    def zeros(self, shape, dtype=torch.float64):
        t = torch.zeros(shape, dtype=dtype)
        self._fields[t.data_ptr()] = ("mid", 0)
        return t

    def register_named(self, t, name):
        self._fields[t.data_ptr()] = (name, 0)

    def peer_ptr(self, t, rank):
        name, _ = self._fields[t.data_ptr()]
        return {"T": 0x100, "T2": 0x200, "mid": 0x300}[name] * 0x100000 + 0x10000000000 * (rank + 1)


def _plan(me, dims, periods, nx=20, ny=24, bank=3):
    fab = MockFabric(me, dims, periods)
    T, T2 = torch.zeros(nx, ny, dtype=torch.float64), torch.zeros(nx, ny, dtype=torch.float64)
    fab.register_named(T, "T")
    fab.register_named(T2, "T2")
    plan = object.__new__(PushPlan)                 # skip the CUDA-only constructor, set what it would set
    plan.fab, plan.bank, plan.epoch, plan.fields, plan.ndim = fab, bank, 0, (T, T2), 2
    plan._desc = {A.data_ptr(): plan._build(A) for A in (T, T2)}
    plan._active, plan._mid, plan._desc2 = True, None, None
    return plan, fab, T, T2


SIDES = [(-1, 0), (1, 0), (0, -1), (0, 1)]


@pytest.mark.parametrize("dims,periods", [((2, 2), (0, 0)), ((3, 2), (0, 0)), ((2, 2), (1, 1)), ((1, 1), (1, 1)),
                                          ((4, 2), (0, 1)), ((1, 3), (1, 0))])
def test_push_and_mid_descriptors_are_wired_symmetrically(dims, periods):
    n = dims[0] * dims[1]
    plans = [_plan(r, dims, periods) for r in range(n)]
    mids = [p.mid_lines() for p, _, _, _ in plans]
    for r, (plan, fab, T, T2) in enumerate(plans):
        d, d2 = plan._desc[T2.data_ptr()], plan._desc2[T2.data_ptr()]
        seen_counters = set()
        for (o, nb) in fab.neighbors:
            k = (o[0] + 1) * 3 + (o[1] + 1)
            kb = (-o[0] + 1) * 3 + (-o[1] + 1)                      # how the neighbour indexes me
            nd = plans[nb][0]._desc[plans[nb][3].data_ptr()]
            assert d.peer_flag[k] == nd.my_flag[kb], "the slot I release is the slot the neighbour polls"
            assert d.peer_dst[k] == fab.peer_ptr(T2, nb)
            assert d2.peer_flag[k] == d.peer_flag[k] and d2.my_flag[k] == d.my_flag[k]      # same DATA flags ...
            assert d2.my_count[k] != d.my_count[k]                                          # ... own counters
            seen_counters |= {d.my_count[k], d2.my_count[k]}
            lo = fab.count_bank_ptr(plan.bank)
            assert lo <= d.my_count[k] < lo + native.BANK_STRIDE * 8 and lo <= d2.my_count[k] < lo + native.BANK_STRIDE * 8
        m = mids[r]
        for s, (ox, oy) in enumerate(SIDES):
            nb = fab.topo.rank_at_offset((ox, oy, 0))
            if nb == PROC_NULL:
                assert not m.peer_mid[s] and not m.peer_flag[s] and not m.my_flag[s]
                continue
            mn = mids[nb]
            assert m.peer_flag[s] == mn.my_flag[s ^ 1], "mid flag: my release slot == the neighbour's poll slot"
            assert m.peer_mid[s] == fab.peer_ptr(plans[r][0]._mid_buf, nb)
            assert m.my_count[s] not in seen_counters                 # mid counters do not collide with tile counters
            pad = fab.bank_ptr(plan.bank)
            assert pad + native.SLOT_MID * 8 <= m.my_flag[s] < pad + native.BANK_STRIDE * 8
            assert m.my_flag[s] != pad + native.SLOT_ERR * 8
        # no mid flag slot coincides with a DATA / READY / ERR slot of the same bank
        used = {d.my_flag[k] for k in range(9) if d.my_flag[k]} | {fab.bank_ptr(plan.bank) + native.SLOT_ERR * 8}
        assert not used & {m.my_flag[s] for s in range(4) if m.my_flag[s]}


def test_two_step_descriptors_need_mid_lines_first():
    plan, fab, T, T2 = _plan(0, (2, 1), (0, 0))
    with pytest.raises(RuntimeError):
        plan.next(T2, two_step=True)
    plan.mid_lines()
    d = plan.next(T2, two_step=True)
    assert d.epoch == 1 and plan.next(T2).epoch == 2          # one epoch sequence for both kinds of launches
