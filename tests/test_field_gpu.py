"""Fq / Fr device arithmetic vs exact integer arithmetic and vs libff (oracle/_ref)."""
import random

import numpy as np
import pytest

import bw6_constants as C
from oracle.ref import ints_to_limbs, limbs_to_ints

pytestmark = pytest.mark.gpu


def _vals(p, n, seed):
    rnd = random.Random(seed)
    edge = [0, 1, 2, p - 1, p - 2, (1 << (p.bit_length() - 1)), (1 << 768) % p, (1 << 384) % p, p >> 1]
    return edge + [rnd.randrange(p) for _ in range(n - len(edge))]


@pytest.mark.parametrize("field", ["fq", "fr"])
def test_field_ops_exact(gpu, field):
    p, n64, rbits = (C.Q, 12, 768) if field == "fq" else (C.R, 6, 384)
    f = gpu.test_fq if field == "fq" else gpu.test_fr
    n = 2048
    a = _vals(p, n, 1)
    b = list(reversed(_vals(p, n, 2)))
    A, B = ints_to_limbs(a, n64), ints_to_limbs(b, n64)
    Rinv = pow(1 << rbits, -1, p)
    assert limbs_to_ints(f(0, A, B)) == [(x + y) % p for x, y in zip(a, b)]
    assert limbs_to_ints(f(1, A, B)) == [(x - y) % p for x, y in zip(a, b)]
    assert limbs_to_ints(f(2, A, B)) == [x * y * Rinv % p for x, y in zip(a, b)]
    if field == "fq":  # op 2 = Karatsuba product + separate reduction (the product path), op 8 = the interleaved CIOS body
        assert np.array_equal(f(8, A, B), f(2, A, B))
        # operands around the 384-bit split of the Karatsuba level: equal halves, halves that differ by one, all-ones halves
        h = (1 << 384) - 1
        s_ = [h, h << 377 & (p - 1), (h << 384 | h) % p, ((h - 1) << 384 | h) % p, (1 << 384) % p, ((1 << 384) + 1) % p, (h << 384) % p]
        sa = [x for x in s_ for _ in s_]
        sb = [y for _ in s_ for y in s_]
        assert limbs_to_ints(f(2, ints_to_limbs(sa, n64), ints_to_limbs(sb, n64))) == [x * y * Rinv % p for x, y in zip(sa, sb)]
    assert limbs_to_ints(f(3, A)) == [x * x * Rinv % p for x in a]
    assert limbs_to_ints(f(5, A)) == [(-x) % p for x in a]
    assert limbs_to_ints(f(6, A)) == [x * Rinv % p for x in a]


@pytest.mark.parametrize("field", ["fq", "fr"])
def test_field_inverse(gpu, field):
    p, n64, rbits = (C.Q, 12, 768) if field == "fq" else (C.R, 6, 384)
    f = gpu.test_fq if field == "fq" else gpu.test_fr
    a = _vals(p, 64, 3)
    A = ints_to_limbs(a, n64)
    Rm = (1 << rbits) % p
    # Montgomery inverse: (aR)^-1 in Montgomery form = a^-1 R = (aR)^-1 R^2
    exp = [0 if x == 0 else pow(x, -1, p) * Rm * Rm % p for x in a]
    assert limbs_to_ints(f(4, A)) == exp
    if field == "fq":  # the Fermat ladder must agree with the binary-Euclid product path
        assert limbs_to_ints(f(7, A)) == exp


def test_field_vs_libff(gpu, ref):
    """Same bytes as libff's Fp_model on the same inputs (fp.tcc:204-227 et al.)."""
    a = ref.fr_rng(100, 512)
    b = ref.fr_rng(900, 512)
    for op in (0, 1, 2, 3, 4, 5, 6):
        assert np.array_equal(gpu.test_fr(op, a, b), ref.fr_op(op, a, b)), op
    rnd = random.Random(5)
    qa = ints_to_limbs([rnd.randrange(C.Q) for _ in range(256)], 12)
    qb = ints_to_limbs([rnd.randrange(C.Q) for _ in range(256)], 12)
    for op in (0, 1, 2, 3, 4, 5, 6):
        assert np.array_equal(gpu.test_fq(op, qa, qb), ref.fq_op(op, qa, qb)), op
