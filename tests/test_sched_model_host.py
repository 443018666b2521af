"""Host model of the score kernel's chunk scheduler (csrc/sched.cuh: static first piece per home CTA, guided
claims on per-row counters, stealing): under random interleavings of the CTAs' atomics every chunk of every row
is posted exactly once, for capped grids, rows of zero units and every chunk size.  The model restates
sched_run's control flow line by line; it guards the claim arithmetic (base offsets of the static pieces), not
the device code's memory ordering, which the GPU tests cover."""
from __future__ import annotations

import random

def simulate(rows_units, grid, groups, cu, min_batch, share_div, seed, tail_chunks=0):
    rnd = random.Random(seed)
    rows = len(rows_units)
    groups = max(1, min(groups, rows))
    share = max(1, grid // groups)
    even = share_div == 0   # the first home rows are cut evenly in units, nothing of them is claimed
    if even:
        even = max(rows_units) * rows * 4 <= sum(rows_units) * 5   # ragged batches keep claims and stealing
        share_div = 1 if even else 2
    counters = [0] * rows
    posted = [[0] * u for u in rows_units]
    chunks_of = lambda r: (rows_units[r] + cu - 1) // cu
    guided = lambda left: max(min_batch, left // (share_div * share))
    static_piece = lambda d: max(1, d // (share_div * share))
    def static_chunks(r):
        if r >= groups or share * groups > grid:
            return 0
        d = chunks_of(r)
        return d if even else min(d, share * static_piece(d))
    def post_units(r, begin, cnt):
        assert 1 <= cnt <= cu
        for u in range(begin, begin + cnt):
            posted[r][u] += 1
    def post(r, c):
        post_units(r, c * cu, min(cu, rows_units[r] - c * cu))
    # each CTA is a generator yielding at every atomic (so interleavings vary)
    def cta(bid):
        def drain(row, is_guided, first_home):
            d = chunks_of(row)
            if d == 0:
                return
            base = static_chunks(row)
            if not is_guided and base + counters[row] >= d:
                return
            if even and base == d:
                if first_home:
                    nu, l = rows_units[row], bid // groups
                    u_end = nu * (l + 1) // share
                    for ub in range(nu * l // share, u_end, cu):
                        post_units(row, ub, min(cu, u_end - ub))
                return
            batch = lambda left: 1 if left <= tail_chunks else (guided(left) if is_guided else min_batch)
            m = static_piece(d) if first_home else batch(d)
            taken = 0
            if first_home:
                j = (bid // groups) * m
                if j >= base:
                    first_home = False
                    m = batch(d)
            if not first_home:
                j = base + counters[row]; counters[row] += m
                yield
            while j < d:
                end = min(j + m, base if first_home else d)
                claimed_to = base if first_home else end
                first_home = False
                taken += end - j
                left = (d - claimed_to) if (is_guided or taken < 12) else 0
                m2 = j2 = 0
                if left > 0:
                    m2 = batch(left)
                    j2 = base + counters[row]; counters[row] += m2
                    yield
                for c in range(j, end):
                    if d - c <= tail_chunks:  # the row's tail: one unit per descriptor
                        for u in range(c * cu, min(c * cu + cu, rows_units[row])):
                            post_units(row, u, 1)
                    else:
                        post(row, c)
                if left <= 0:
                    return
                j, m = j2, m2
        if bid < share * groups:
            grp = bid % groups
            for row in range(grp, rows, groups):
                yield from drain(row, True, row == grp)
        while True:
            best, best_row = 0, -1
            for r in range(rows):
                rr = (r + bid) % rows
                left = chunks_of(rr) - static_chunks(rr) - counters[rr]
                if left > best or (left == best and left > 0 and rr > best_row):
                    best, best_row = left, rr
            yield
            if best <= 0:
                break
            yield from drain(best_row, False, False)
    live = [cta(b) for b in range(grid)]
    while live:
        g = rnd.choice(live)
        try:
            next(g)
        except StopIteration:
            live.remove(g)
    for r in range(rows):
        assert all(x == 1 for x in posted[r]), (rows_units, grid, groups, cu, min_batch, r, posted[r][:20])



def test_every_unit_is_posted_exactly_once():
    rnd = random.Random(0)
    for trial in range(600):
        rows = rnd.choice([1, 2, 3, 5, 16, 32])
        units = [rnd.choice([0, 1, 2, 7, 33, 128, 512]) for _ in range(rows)] if trial % 3 else [rnd.choice([1, 7, 100, 512])] * rows
        grid = rnd.choice([1, 2, 7, 16, 37, 128, 148])
        groups = rnd.choice([rows, max(1, rows // 2), 1])
        simulate(units, grid, groups, rnd.choice([1, 2, 3, 4, 8]), rnd.choice([1, 2, 3]), rnd.choice([0, 1, 2, 4]), trial,
                 tail_chunks=rnd.choice([0, 0, 1, 5, 40]))
