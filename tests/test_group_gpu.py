"""G1/G2 device point arithmetic vs libff (bw6_761_g1.cpp / bw6_761_g2.cpp), normalised results."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("group", [1, 2])
def test_group_ops(gpu, ref, group):
    n = 64
    s1 = ref.fr_rng(10, n)
    s2 = ref.fr_rng(20, n)
    P = ref.bases_mul(s1, group)   # special form
    Q = ref.bases_mul(s2, group)
    zero = ref.group_op(1, P[:1], P[:1], group) * 0
    # craft the special cases: P+P, P+(-P), 0+Q, P+0
    Qs = Q.copy()
    Ps = P.copy()
    Qs[0] = P[0]
    Qs[1] = ref.group_op(4, P[1:2], None, group)[0]
    inf = np.zeros(36, dtype=np.uint64)
    inf[12:24] = P[0][24:36]  # (0, one, 0)
    Ps[2] = inf
    Qs[3] = inf
    Ps[4] = inf
    Qs[4] = inf
    norm = lambda x: ref.group_op(3, x, None, group)
    assert np.array_equal(gpu.test_group(0, Ps, Qs), norm(ref.group_op(0, Ps, Qs, group)))
    assert np.array_equal(gpu.test_group(1, Ps, Qs), norm(ref.group_op(1, Ps, Qs, group)))
    assert np.array_equal(gpu.test_group(2, Ps), norm(ref.group_op(2, Ps, None, group)))
    # non-normalised projective inputs: (P+Q) computed by libff, then doubled / added on device
    S = ref.group_op(0, P, Q, group)
    assert np.array_equal(gpu.test_group(2, S), norm(ref.group_op(2, S, None, group)))
    assert np.array_equal(gpu.test_group(0, S, ref.group_op(2, Q, None, group)),
                          norm(ref.group_op(0, S, ref.group_op(2, Q, None, group), group)))
    assert np.array_equal(gpu.test_group(3, S), norm(S))


def _synth_scalar(seed, i):
    M = (1 << 64) - 1
    x = (seed * 0x9E3779B97F4A7C15 + i * 0xBF58476D1CE4E5B9 + 0x94D049BB133111EB) & M
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & M
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & M
    x ^= x >> 31
    return (x & 0xFFFFFFFF) or 1


@pytest.mark.parametrize("group", [1, 2])
def test_synth_bases_are_multiples_of_the_generator(gpu, ref, group):
    n = 300
    got = gpu.synth_bases(n, group, seed=5)
    ks = ref.fr_from_ints([_synth_scalar(5, i) for i in range(n)])
    assert np.array_equal(got, ref.bases_mul(ks, group))
