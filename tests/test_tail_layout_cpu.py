"""Index arithmetic of the MSM tail (msm.cuh: k_tail_level / k_tail_scale / k_tail_sum) replayed on the
integers: the group is (Z, +), a "point" is a Python int, doubling is *2. The model below follows the
kernels line by line (same buffers, same offsets, one thread = one addition), so a wrong offset or a
clobbered ping-pong buffer shows up here without a GPU. Checked against the definition
    result = sum_v 2^(c v) * sum_b (b + 1) * B[v][b]          (multiexp.tcc:90-125 and :614-616)."""
import random

import pytest


def tail_level_threads(cm1, V, k):
    if k > cm1:
        return 0
    nk = (1 << cm1) >> k
    return V * nk + (k * V * (nk >> 1) if k < cm1 else 0)


def run_tail(buckets, c, V):
    """buckets[v*NB + b] = bucket sum or None (empty)."""
    cm1 = c - 1
    NB = 1 << cm1
    size = V * NB - (V * NB) // 4
    lvl = [[None] * size, [None] * size]
    launches = 0
    for k in range(1, cm1 + 1):
        nk = NB >> k
        n_main, n_tree = V * nk, V * (nk >> 1)
        src, out = lvl[(k - 1) & 1], lvl[k & 1]
        writes = {}

        def x_prev(i):
            return (buckets[i] or 0) if k == 1 else src[i]

        for tid in range(tail_level_threads(cm1, V, k)):
            if tid < n_main:
                a, b, dst = x_prev(2 * tid), x_prev(2 * tid + 1), tid
            else:
                t = tid - n_main
                tj, u = divmod(t, n_tree)
                if tj == k - 1:
                    a, b = x_prev(4 * u + 1), x_prev(4 * u + 3)
                else:
                    base = 2 * n_main + tj * (2 * n_tree)
                    a, b = src[base + 2 * u], src[base + 2 * u + 1]
                dst = n_main + tj * n_tree + u
            assert a is not None and b is not None, (k, tid)
            assert dst < size and dst not in writes
            writes[dst] = a + b
        for d, val in writes.items():  # all threads of a launch read before anyone's write is visible
            out[d] = val
        launches += 1
    # k_tail_scale
    fin = [None] * (V * c)
    s2 = lvl[(cm1 - 1) & 1]
    for tid in range(V * c):
        v, j = divmod(tid, c)
        ndbl = c * v
        if j == cm1:
            srcv = lvl[cm1 & 1][v]
        else:
            srcv = s2[2 * v + 1] if j + 1 == cm1 else s2[2 * V + j * V + v]
            ndbl += j
        fin[tid] = srcv << ndbl
    # k_tail_sum: pair tree into fin[0]
    n, step = V * c, 1
    while step < n:
        n_pairs = (n - step + 2 * step - 1) // (2 * step)
        for q in range(n_pairs):
            fin[2 * step * q] += fin[2 * step * q + step]
        step <<= 1
    return fin[0], launches


@pytest.mark.parametrize("c,V", [(4, 1), (4, 3), (5, 1), (6, 2), (8, 1), (9, 5), (11, 1)])
def test_tail_dag_matches_definition(c, V):
    rng = random.Random(c * 100 + V)
    NB = 1 << (c - 1)
    buckets = [rng.randrange(1, 1 << 40) if rng.random() < 0.8 else None for _ in range(V * NB)]
    expect = sum(((b + 1) * (buckets[v * NB + b] or 0)) << (c * v) for v in range(V) for b in range(NB))
    got, launches = run_tail(buckets, c, V)
    assert got == expect
    assert launches == c - 1


def test_tail_thread_counts_sum_to_two_nb():
    """Work of the DAG = the running sum's: ~2 NB additions per bucket set."""
    for c in (8, 16, 18):
        cm1 = c - 1
        total = sum(tail_level_threads(cm1, 1, k) for k in range(1, cm1 + 1))
        assert (1 << cm1) * 2 - 2 * c <= total <= (1 << cm1) * 2
