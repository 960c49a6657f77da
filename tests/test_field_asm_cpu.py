"""The generated PTX carry-chain bodies (tools/gen_field_asm.py -> csrc/field_asm.cuh) run through a tiny PTX
interpreter (tools/ptx_sim.py) and compared with exact integers: Montgomery product and the reduced
squaring for Fq (24 limbs) and Fr (12 limbs). Reference semantics: Fp_model::mul_reduce / squared,
libff/algebra/fields/fp.tcc:204-227, 597-612 (canonical representative in [0, p))."""
import random

import pytest

import bw6_constants as C
import gen_field_asm as G
import ptx_sim


def _run(em, N, a, b=None):
    regs = {}
    for k in range(N):
        regs["%%%d" % (N + k)] = (a >> (32 * k)) & 0xFFFFFFFF
        if b is not None:
            regs["%%%d" % (2 * N + k)] = (b >> (32 * k)) & 0xFFFFFFFF
    ptx_sim.run(em.lines, regs)
    return sum(regs["%%%d" % k] << (32 * k) for k in range(N))


def _cases(p, n, seed):
    rnd = random.Random(seed)
    top = (1 << p.bit_length()) - 1
    edge = [0, 1, 2, p - 1, p - 2, (p - 1) // 2, (p + 1) // 2, top & (p - 1), (1 << (p.bit_length() - 1)),
            int("ffffffff" * 40, 16) % p, int("80000000" * 40, 16) % p, int("7fffffff" * 40, 16) % p,
            int("ffffffff00000000" * 20, 16) % p, int("00000000ffffffff" * 20, 16) % p]
    return edge + [rnd.randrange(p) for _ in range(n)]


@pytest.mark.parametrize("tag,N,P,K", [("fq", C.FQ_LIMBS32, C.Q, C.FQ), ("fr", C.FR_LIMBS32, C.R, C.FR)])
def test_generated_mul_and_sqr(tag, N, P, K):
    r_op = lambda k: "%%%d" % k
    a_op = lambda k: "%%%d" % (N + k)
    b_op = lambda k: "%%%d" % (2 * N + k)
    mul = G.gen_mul(N, P, K["inv32"], a_op, b_op, r_op)
    sqr = G.gen_mul(N, P, K["inv32"], a_op, a_op, r_op, square=True)
    rinv = pow(1 << (32 * N), -1, P)
    n_wide = lambda em: sum(1 for l in em.lines if l.startswith(("mul.lo", "mad.lo", "madc.lo")))
    assert n_wide(mul) == 2 * N * N + N                       # + the N single-word m = T0 * inv
    assert n_wide(sqr) == 2 * N * N - N * N // 4 + N   # even columns below the diagonal dropped
    xs = _cases(P, 60 if N == 24 else 150, 5)
    for i, a in enumerate(xs):
        assert _run(sqr, N, a) == a * a * rinv % P, hex(a)
        b = xs[(7 * i + 3) % len(xs)]
        assert _run(mul, N, a, b) == a * b * rinv % P
