"""Host model of the attention kernel's cost-aware stream-K cut (csrc/decode_attn.cu, struct AttnCut).

The kernel cuts a virtual sequence (the rows' 16-token units with `pen` padding positions at every row boundary)
into G equal ranges.  This model restates that arithmetic in Python and checks the invariants the kernel's epilogue
relies on, for ragged rows, every grid size and every penalty the launcher can pass:
  * the CTAs' unit ranges partition [0, U) in order;
  * owner(u, r) is the CTA whose range holds unit u;
  * the contributors of a row are exactly owner(first unit) .. owner(last unit) and none of them is empty
    (the arrival ticket counts on it);
  * a CTA has at most one incomplete segment at either end (two partial slots per CTA are enough).
"""
import random

import pytest


class Cut:
    def __init__(self, units, grid, split_pen):
        self.prefix = [0]
        for n in units:
            self.prefix.append(self.prefix[-1] + n)
        self.R = len(units)
        self.U = self.prefix[-1]
        self.G = min(self.U, grid)
        self.pen = min(split_pen, self.U // self.G // 2) if self.G > 0 else 0
        self.UV = self.U + self.pen * (self.R - 1)

    def vbegin(self, r):
        return self.prefix[r] + self.pen * r

    def row_of_virtual(self, v):
        lo, hi = 0, self.R
        while hi - lo > 1:
            mid = (lo + hi) >> 1
            if self.vbegin(mid) <= v:
                lo = mid
            else:
                hi = mid
        return lo

    def to_unit(self, v):
        r = self.row_of_virtual(v)
        return self.prefix[r] + min(v - self.vbegin(r), self.prefix[r + 1] - self.prefix[r])

    def begin_of(self, c):
        return self.to_unit(self.UV * c // self.G)

    def range_of(self, c):
        return self.begin_of(c), (self.U if c + 1 == self.G else self.begin_of(c + 1))

    def owner(self, u, r):
        v = u + self.pen * r
        return ((v + 1) * self.G + self.UV - 1) // self.UV - 1


def check(units, grid, split_pen):
    cut = Cut(units, grid, split_pen)
    if cut.U == 0:
        return cut
    ranges = [cut.range_of(c) for c in range(cut.G)]
    assert ranges[0][0] == 0 and ranges[-1][1] == cut.U
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a0 <= a1
    holder = {}
    for c, (u0, u1) in enumerate(ranges):
        for u in range(u0, u1):
            holder[u] = c
    for r in range(cut.R):
        b, e = cut.prefix[r], cut.prefix[r + 1]
        if b == e:
            continue
        for u in (b, e - 1, (b + e) // 2):
            assert cut.owner(u, r) == holder[u], (units, grid, split_pen, r, u)
        first, last = cut.owner(b, r), cut.owner(e - 1, r)
        contributors = sorted({holder[u] for u in range(b, e)})
        assert contributors == list(range(first, last + 1)), (units, grid, split_pen, r)
    for c, (u0, u1) in enumerate(ranges):
        incomplete = 0
        for r in range(cut.R):
            b, e = cut.prefix[r], cut.prefix[r + 1]
            s0, s1 = max(b, u0), min(e, u1)
            if s0 < s1 and not (s0 == b and s1 == e):
                incomplete += 1
                assert s0 == u0 or s1 == u1
        assert incomplete <= 2
    return cut


def test_uniform_rows_cfg5_shape():
    cut = check([257] * 32, 148, 8)
    sizes = [b - a for a, b in (cut.range_of(c) for c in range(cut.G))]
    split = [c for c in range(cut.G)
             if any(cut.range_of(c)[0] < cut.prefix[r] < cut.range_of(c)[1] for r in range(1, cut.R))]
    # a CTA with a row boundary inside streams pen units less than a CTA in the middle of a row (a CTA whose range
    # ENDS at a boundary may lose part of the padding too: it is not split, just shorter)
    full = max(sizes)
    assert split and all(full - cut.pen - 2 <= sizes[c] <= full - cut.pen + 1 for c in split)
    assert len(split) <= cut.R - 1 and sum(sizes) == cut.U


@pytest.mark.parametrize("seed", range(6))
def test_ragged_rows_random(seed):
    rng = random.Random(seed)
    for _ in range(150):
        rows = rng.randint(1, 40)
        units = [rng.choice([0, 1, 2, 3, 7, 33, 129, 257, 600]) if rng.random() < 0.7 else rng.randint(0, 300)
                 for _ in range(rows)]
        check(units, rng.choice([1, 2, 3, 7, 32, 131, 148]), rng.choice([0, 1, 4, 8, 16, 32, 64]))


def test_penalty_zero_is_the_equal_cut():
    cut = check([100, 50, 7, 300], 148, 0)
    assert all(cut.range_of(c) == (cut.U * c // cut.G, cut.U * (c + 1) // cut.G) for c in range(cut.G))
