"""Host restatement of the slot arithmetic of the batched-affine pair tree (k_tree_round / k_tree_finish in
zecale_b200/csrc/msm.cuh): level r of bucket b starts at (off[b] >> r) + b, a round maps output slot s to
(bucket, pair) by a search on that closed form, and after ceil(log2(max count)) rounds every non-empty bucket
holds the sum of its entries. Points are replaced by integers (the group is abelian, so only the pairing matters)."""
import numpy as np
import pytest


def _base(off, b, r):
    return (int(off[b]) >> r) + b


def _round(r, off, nb, level_in, entries):
    """One k_tree_round: returns the level r+1 array (None = slot never written)."""
    M = int(off[nb])
    slots = (M >> (r + 1)) + nb
    out = [None] * slots
    for s in range(slots):
        lo, hi = 0, nb - 1                      # largest b with base_{r+1}(b) <= s
        while lo < hi:
            mid = (lo + hi + 1) >> 1
            if _base(off, mid, r + 1) <= s:
                lo = mid
            else:
                hi = mid - 1
        b = lo
        i = s - _base(off, b, r + 1)
        n_b = int(off[b + 1] - off[b])
        cr = (n_b + (1 << r) - 1) >> r
        if i >= (cr + 1) >> 1:
            continue                            # padding slot
        pos = (int(off[b]) if r == 0 else _base(off, b, r)) + 2 * i
        src = entries if r == 0 else level_in
        out[s] = src[pos] if 2 * i + 1 >= cr else src[pos] + src[pos + 1]
    return out


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_pair_tree_layout_and_sums(seed):
    rng = np.random.default_rng(seed)
    nb = 37
    counts = rng.integers(0, 9, size=nb)
    counts[rng.integers(0, nb)] = 150           # a heavy bucket decides the number of rounds
    counts[0] = 0 if seed % 2 else 1
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
    entries = [int(v) for v in rng.integers(1, 1 << 40, size=int(off[nb]))]
    # buckets never overlap at any level
    for r in range(1, 10):
        for b in range(nb):
            cr = (int(counts[b]) + (1 << r) - 1) >> r
            assert _base(off, b, r) + cr <= _base(off, b + 1, r)
    rounds = 0
    while (1 << rounds) < int(counts.max()):
        rounds += 1
    level = None
    for r in range(rounds):
        level = _round(r, off, nb, level, entries)
    for b in range(nb):
        if counts[b] == 0:
            continue
        got = entries[int(off[b])] if rounds == 0 else level[_base(off, b, rounds)]
        assert got == sum(entries[int(off[b]):int(off[b + 1])])
