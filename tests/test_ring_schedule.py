"""The per-thread landmark ring of k_fs2_update_fast (slam_b200/csrc/kernels.cuh): a model of its copy schedule.

The kernel keeps D = SLAM_RING_DEPTH rows in flight per thread with cp.async: one commit group per row (an empty group
where there is nothing to fetch), `cp.async.wait_group D-1` before a slot is read.  Pass 1 walks the observations
forwards, pass 2 backwards starting from the rows pass 1 left in the ring.  The KFULL instantiation (K a multiple of D)
replaces the per-row "is there a row to prefetch" tests by one test per D rows.  This model replays both variants'
control flow for every K and checks what the kernel relies on: a slot holds the row that is read from it, that row's
group has completed by the wait in front of the read, no slot is overwritten before its last read, and every row is
fetched exactly once per pass.  (A restatement of the schedule, not the kernel: the GPU parity tests cover the kernel.)"""
import pytest

D = 4


class Ring:
    def __init__(self):
        self.slot = [None] * D      # row whose copy was last issued into the slot
        self.group_of = {}          # (pass, row) -> index of the commit group its copy belongs to
        self.groups = 0             # commit groups so far
        self.pending = []           # copies issued since the last commit
        self.fetches = []

    def fetch(self, tag, row, slot):
        self.slot[slot] = (tag, row)
        self.pending.append((tag, row))
        self.fetches.append((tag, row))

    def commit(self):
        for c in self.pending:
            self.group_of[c] = self.groups
        self.pending = []
        self.groups += 1

    def landed(self, tag, row, n_outstanding):
        """wait_group n: all but the newest n groups have completed."""
        return self.group_of[(tag, row)] < self.groups - n_outstanding


def run(K, kfull):
    r = Ring()
    for j in range(D - 1):                       # prologue
        if kfull or j < K:
            r.fetch(1, j, j)
        r.commit()
    i0 = 0
    while i0 < K:                                # pass 1, forwards
        more = i0 + D < K
        for sl in range(D):
            i = i0 + sl
            if kfull or i < K:
                j = i + D - 1
                if ((sl == 0 or more) if kfull else (j < K)):
                    assert j < K, "prefetch of a row that does not exist"
                    r.fetch(1, j, (sl + D - 1) % D)
                r.commit()
                assert r.slot[sl] == (1, i), (K, kfull, "pass 1 reads", i, "slot holds", r.slot[sl])
                assert r.landed(1, i, D - 1)
        i0 += D
    assert sorted(x for t, x in r.fetches if t == 1) == list(range(K))
    i0 = ((K - 1) // D) * D
    while i0 >= 0:                               # pass 2, backwards
        more = i0 >= D
        for sl in range(D - 1, -1, -1):
            i = i0 + sl
            if kfull or i < K:
                held = r.slot[sl]
                assert held[1] == i, (K, kfull, "pass 2 reads", i, "slot holds", held)
                assert r.landed(held[0], i, D - 1)
                if more:
                    r.fetch(2, i - D, sl)        # the slot just read takes row i - D
                r.commit()
        i0 -= D
    # pass 2 re-fetches exactly the rows pass 1 did not leave in the ring (the last D rows stay), each once
    assert sorted(x for t, x in r.fetches if t == 2) == list(range(max(0, K - D)))
    return r


@pytest.mark.parametrize("K", range(1, 70))
def test_ragged_schedule(K):
    run(K, kfull=False)


@pytest.mark.parametrize("K", range(D, 70, D))
def test_kfull_schedule_equals_the_ragged_one(K):
    a, b = run(K, kfull=True), run(K, kfull=False)
    assert a.fetches == b.fetches and a.groups == b.groups
