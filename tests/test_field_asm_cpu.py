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


@pytest.mark.parametrize("levels", [1, 2])
def test_generated_karatsuba_product(levels):
    """Round 2 product: Karatsuba on a*b, word-serial reduction on the low half, P_hi added at the end
    (gen_mul_kara) — the same canonical value as the interleaved form."""
    N, P, K = C.FQ_LIMBS32, C.Q, C.FQ
    r_op = lambda k: "%%%d" % k
    a_op = lambda k: "%%%d" % (N + k)
    b_op = lambda k: "%%%d" % (2 * N + k)
    em = G.gen_mul_kara(N, P, K["inv32"], a_op, b_op, r_op, levels)
    assert G.n_wide(em) == {1: 3 * 144, 2: 9 * 36}[levels] + N * N + N
    rinv = pow(1 << (32 * N), -1, P)
    xs = _cases(P, 25, 11) + [(1 << 384) - 1, 1 << 384, ((1 << 377) - 1) << 384, ((1 << 384) - 1) << 370 & ((1 << 761) - 1)]
    xs = [x % P for x in xs]
    for i, a in enumerate(xs):
        for b in (xs[(7 * i + 3) % len(xs)], a):
            assert _run(em, N, a, b) == a * b * rinv % P, (hex(a), hex(b))


def test_generated_wide_product_difference():
    """fq_mulw_regs / fq_wsubk_regs / fq_redc_regs (the VM's MULW + MULFIN): (a*b - c*e) R^-1 mod q with one reduction."""
    N, P, K = C.FQ_LIMBS32, C.Q, C.FQ
    pc = lambda base: (lambda k: "%%%d" % (base + k))
    mulw = G.gen_mulw(N, pc(2 * N), pc(3 * N), pc(0), 1)
    wsubk = G.gen_wsubk(N, P, pc(0), pc(2 * N))
    redc = G.gen_redc_only(N, P, K["inv32"], pc(N), pc(0))
    rinv = pow(1 << (32 * N), -1, P)

    def put(x, n, base):
        return {"%%%d" % (base + k): (x >> (32 * k)) & 0xFFFFFFFF for k in range(n)}

    def get(regs, n, base):
        return sum(regs["%%%d" % (base + k)] << (32 * k) for k in range(n))

    def do(em, outs, **ins):
        regs = {}
        for (x, n, base) in ins.values():
            regs.update(put(x, n, base))
        ptx_sim.run(em.lines, regs)
        return get(regs, *outs)

    xs = _cases(P, 12, 13)
    for i in range(len(xs)):
        a, b, c, e = xs[i], xs[(3 * i + 1) % len(xs)], xs[(5 * i + 2) % len(xs)], xs[(7 * i + 3) % len(xs)]
        w = do(mulw, (2 * N, 0), a=(a, N, 2 * N), b=(b, N, 3 * N))
        assert w == a * b
        t = do(mulw, (2 * N, 0), a=(c, N, 2 * N), b=(e, N, 3 * N))
        d = do(wsubk, (2 * N, 0), w=(w, 2 * N, 0), p=(t, 2 * N, 2 * N))
        assert d == a * b - c * e + P * P
        assert do(redc, (N, 0), p=(d, 2 * N, N)) == (a * b - c * e) * rinv % P
    for w, t in ((0, (P - 1) ** 2), ((P - 1) ** 2, 0), (0, 0)):
        d = do(wsubk, (2 * N, 0), w=(w, 2 * N, 0), p=(t, 2 * N, 2 * N))
        assert do(redc, (N, 0), p=(d, 2 * N, N)) == (w - t) * rinv % P


def test_generated_fr_product_with_inv_operand():
    """The shipped Fr body takes inv = -r^-1 mod 2^32 as an operand (through the constant bank on the device) instead of as
    an immediate; same values."""
    N, P, K = C.FR_LIMBS32, C.R, C.FR
    r_op = lambda k: "%%%d" % k
    a_op = lambda k: "%%%d" % (N + k)
    b_op = lambda k: "%%%d" % (2 * N + k)
    em = G.gen_mul(N, P, K["inv32"], a_op, b_op, r_op, q_op=lambda k: "%%%d" % (3 * N), inv_only=True)
    assert G.n_wide(em) == 2 * N * N + N
    rinv = pow(1 << (32 * N), -1, P)
    xs = _cases(P, 60, 17)
    for i, a in enumerate(xs):
        b = xs[(5 * i + 1) % len(xs)]
        regs = {"%%%d" % (3 * N): K["inv32"]}
        for k in range(N):
            regs["%%%d" % (N + k)] = (a >> (32 * k)) & 0xFFFFFFFF
            regs["%%%d" % (2 * N + k)] = (b >> (32 * k)) & 0xFFFFFFFF
        ptx_sim.run(em.lines, regs)
        assert sum(regs["%%%d" % k] << (32 * k) for k in range(N)) == a * b * rinv % P
