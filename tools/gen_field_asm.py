#!/usr/bin/env python3
"""Generate zecale_b200/csrc/field_asm.cuh: PTX carry-chain bodies for BW6-761 Fq/Fr.

Montgomery multiplication (word-serial CIOS, 32-bit limbs, R = 2^(32 N)):

The running value T is kept as two limb vectors E ("even-aligned") and O
("odd-aligned", weight 2^32) with T = E + O*2^32 + x (x = one pending word at
column 0).  For one word b_i of b the products a_j*b_i with even j land on limb
pairs (E_j, E_j+1) and those with odd j on (O_j-1, O_j), so each half is ONE
uninterrupted carry chain of mad.lo.cc / madc.hi.cc pairs (ptxas fuses a pair
into one IMAD.WIDE.U32 with a predicate carry).  After adding m*q (m = E_0*inv)
the low word is zero and the division by 2^32 is pure renaming:
    E' = O (+ E_1 at column 0, carry folded into the next odd chain),
    O' = E >> 64.
No limb ever ripples; per word of b the cost is 2N wide MADs + 1 MUL + 3 ADD.
Bounds (both moduli leave 7 spare bits in the top limb): E < 2^(32N) on entry,
top carry word <= 2, the odd chain never carries out.

The same split is what libff's Fp_model::mul_reduce computes (fp.tcc:204-227:
mpn_mul_n then N x mpn_addmul_1 + final conditional subtract); results are the
canonical representative in [0, p) so they are bit-identical.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(__file__))
import bw6_constants as C


class Emit:
    def __init__(self):
        self.lines = []
        self.nreg = 0

    def fresh(self):
        self.nreg += 1
        return "v%d" % (self.nreg - 1)

    def op(self, s):
        self.lines.append(s)


def imm(x):
    return "0x%08x" % x


def gen_mul(N, P, inv, a_op, b_op, r_op, square=False, q_op=None, inv_only=False):
    """a_op(j), b_op(i), r_op(k) -> PTX operand strings (%k placeholders). q_op(j): operand holding limb j of the
    modulus in the m*q chains (default: an immediate)."""
    q = C.limbs32(P, N)
    em = Emit()
    qm = (lambda j: q_op(j)) if (q_op and not inv_only) else (lambda j: imm(q[j]))
    invm = q_op(N) if q_op else imm(inv)          # with q_op, operand N of the modulus array holds inv
    E = [em.fresh() for _ in range(N)]
    O = [em.fresh() for _ in range(N)]
    top = em.fresh()
    x = None
    if square:
        # a^2 = sum_r a_r X_r 2^(32 r). The pair (i, j), j < i, is owed 2 a_i a_j:
        #   j even: row j carries it doubled (column i), row i drops column j;
        #   j odd : rows i and j each carry it once.
        # So an even row is  [odd columns below r: a_c | a_r | (a_(r+1) << 1) | limbs of 2a above],
        # an odd row is a with the even columns below r dropped, and only EVEN columns below the
        # diagonal disappear: the even carry chain just starts later, the odd chain stays whole (the
        # carry of the pending-word fold enters it), no carry has to ripple through empty columns.
        # Saves N^2/4 wide MADs of the N^2 in the multiply half (144 of 576 for Fq).
        D = {}
        for j in range(2, N):
            D[j] = em.fresh()
            em.op("shf.l.wrap.b32 %s, %s, %s, 1;" % (D[j], a_op(j - 1), a_op(j)))
    for i in range(N):
        bi = b_op(i)
        m = em.fresh()
        if square:
            row = {c: a_op(c) for c in range(N) if not (c < i and c % 2 == 0)}
            if i % 2 == 0:
                if i + 1 < N:
                    row[i + 1] = em.fresh()
                    em.op("shl.b32 %s, %s, 1;" % (row[i + 1], a_op(i + 1)))
                for j in range(i + 2, N):
                    row[j] = D[j]
            col = lambda j, row=row: row.get(j)
        else:
            col = a_op
        if i == 0:
            for j in range(0, N, 2):
                em.op("mul.lo.u32 %s, %s, %s;" % (E[j], col(j), bi))
                em.op("mul.hi.u32 %s, %s, %s;" % (E[j + 1], col(j), bi))
            for j in range(1, N, 2):
                em.op("mul.lo.u32 %s, %s, %s;" % (O[j - 1], col(j), bi))
                em.op("mul.hi.u32 %s, %s, %s;" % (O[j], col(j), bi))
            em.op("mul.lo.u32 %s, %s, %s;" % (m, E[0], invm))
            # even q chain
            for j in range(0, N, 2):
                c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c0, E[j], m, qm(j), E[j]))
                em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], m, qm(j), E[j + 1]))
            em.op("addc.u32 %s, 0, 0;" % top)
        else:
            # fold pending word; its carry (weight 2^32) enters the odd chain
            em.op("add.cc.u32 %s, %s, %s;" % (E[0], E[0], x))
            for j in range(1, N, 2):
                if col(j) is None:  # no product in this column (squaring): the carry of the fold just ripples
                    em.op("addc.cc.u32 %s, %s, 0;" % (O[j - 1], O[j - 1]))
                    em.op("addc.cc.u32 %s, %s, 0;" % (O[j], O[j]))
                    continue
                em.op("madc.lo.cc.u32 %s, %s, %s, %s;" % (O[j - 1], col(j), bi, O[j - 1]))
                if j == N - 1:
                    nw = em.fresh()
                    em.op("madc.hi.u32 %s, %s, %s, 0;" % (nw, col(j), bi))
                    O[j] = nw
                else:
                    em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (O[j], col(j), bi, O[j]))
            first = True
            for j in range(0, N, 2):
                if col(j) is None:  # below the diagonal: nothing to add and no carry yet
                    continue
                c0 = "mad.lo.cc.u32" if first else "madc.lo.cc.u32"
                first = False
                em.op("%s %s, %s, %s, %s;" % (c0, E[j], col(j), bi, E[j]))
                em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], col(j), bi, E[j + 1]))
            top = em.fresh()
            if first:
                em.op("add.u32 %s, 0, 0;" % top)
            else:
                em.op("addc.u32 %s, 0, 0;" % top)
            em.op("mul.lo.u32 %s, %s, %s;" % (m, E[0], invm))
            for j in range(0, N, 2):
                c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c0, E[j], m, qm(j), E[j]))
                em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], m, qm(j), E[j + 1]))
            em.op("addc.u32 %s, %s, 0;" % (top, top))
        # odd q chain (never carries out)
        for j in range(1, N, 2):
            c0 = "mad.lo.cc.u32" if j == 1 else "madc.lo.cc.u32"
            em.op("%s %s, %s, %s, %s;" % (c0, O[j - 1], m, qm(j), O[j - 1]))
            c1 = "madc.hi.u32" if j == N - 1 else "madc.hi.cc.u32"
            em.op("%s %s, %s, %s, %s;" % (c1, O[j], m, qm(j), O[j]))
        # divide by 2^32: rename
        x = E[1]
        newO = E[2:] + [top, None]
        E = O
        O = newO
    # merge: T = E + x + O*2^32   (O[N-1] is None == 0)
    em.op("add.cc.u32 %s, %s, %s;" % (E[0], E[0], x))
    for k in range(1, N):
        cc = "addc.cc.u32" if k < N - 1 else "addc.u32"
        em.op("%s %s, %s, %s;" % (cc, E[k], E[k], O[k - 1]))
    gen_final_sub(em, N, q, E, r_op)
    return em


def gen_final_sub(em, N, q, T, r_op):
    """r = T - q if T >= q else T   (T < 2q, N limbs)."""
    D = [em.fresh() for _ in range(N)]
    for k in range(N):
        cc = "sub.cc.u32" if k == 0 else "subc.cc.u32"
        em.op("%s %s, %s, %s;" % (cc, D[k], T[k], imm(q[k])))
    bw = em.fresh()
    em.op("subc.u32 %s, 0, 0;" % bw)  # 0xffffffff if T < q
    em.op("setp.ne.u32 pb, %s, 0;" % bw)
    for k in range(N):
        em.op("selp.u32 %s, %s, %s, pb;" % (r_op(k), T[k], D[k]))


def gen_add(N, P, a_op, b_op, r_op):
    q = C.limbs32(P, N)
    em = Emit()
    T = [em.fresh() for _ in range(N)]
    for k in range(N):
        cc = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < N - 1 else "addc.u32")
        em.op("%s %s, %s, %s;" % (cc, T[k], a_op(k), b_op(k)))
    gen_final_sub(em, N, q, T, r_op)
    return em


def gen_sub(N, P, a_op, b_op, r_op):
    """r = a - b (+ q if borrow)."""
    q = C.limbs32(P, N)
    em = Emit()
    T = [em.fresh() for _ in range(N)]
    for k in range(N):
        cc = "sub.cc.u32" if k == 0 else "subc.cc.u32"
        em.op("%s %s, %s, %s;" % (cc, T[k], a_op(k), b_op(k)))
    bw = em.fresh()
    em.op("subc.u32 %s, 0, 0;" % bw)  # all ones if a < b
    # masked add of q
    for k in range(N):
        mk = em.fresh()
        em.op("and.b32 %s, %s, %s;" % (mk, bw, imm(q[k])))
        cc = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < N - 1 else "addc.u32")
        em.op("%s %s, %s, %s;" % (cc, r_op(k), T[k], mk))
    return em


# ---------------------------------------------------------------------------------------------
# Round 2: product and reduction separated, Karatsuba on the product.
#
# r = (P + M*q)/R = P_hi + (P_lo + M*q)/R,  P = a*b (2N limbs),  M = sum m_i 2^(32 i) chosen so that
# P_lo + M*q = 0 mod R. The m_i depend on P_lo and q only, so the word-serial rounds run on an N-limb
# window that starts as P_lo (T <- (T + m*q) >> 32, m = T_0 * inv: the m*q half of the CIOS round above,
# N wide MADs instead of 2N) and P_hi is added once at the end: T_N <= q, P_hi < q^2/R < q, sum < 2q,
# one conditional subtract -- the same canonical value. The product a*b no longer has to be interleaved
# with the reduction, so it can be Karatsuba: a = a0 + a1 B, b = b0 + b1 B,
#     a*b = z0 + (z0 + z2 - (a0 - a1)(b0 - b1)) B + z2 B^2,   z0 = a0 b0, z2 = a1 b1,
# three half-size products instead of four; (a0 - a1)(b0 - b1) = +-|a0 - a1| |b0 - b1| with the sign
# applied as a masked two's complement. Each level trades N^2/4 wide MADs (the saturated pipe,
# DESIGN.md section 3.1) for ~8N adds/logic ops on the otherwise idle ALU pipe. Base case: schoolbook rows
# in the even/odd carry-chain form of gen_mul.
#   Fq: 1152 + 24 wide MADs  ->  levels=1: 432 + 576 + 24,  levels=2: 324 + 576 + 24.


def school_product(em, A, B):
    """A, B: n operand names (n even). Returns 2n names holding A*B. n*n wide MADs."""
    n = len(A)
    assert n == len(B) and n % 2 == 0
    E = [em.fresh() for _ in range(n)]
    O = [em.fresh() for _ in range(n)]
    out = []
    x = None
    tops = {}
    for i in range(n):
        bi = B[i]
        if i == 0:
            for j in range(0, n, 2):
                em.op("mul.lo.u32 %s, %s, %s;" % (E[j], A[j], bi))
                em.op("mul.hi.u32 %s, %s, %s;" % (E[j + 1], A[j], bi))
            for j in range(1, n, 2):
                em.op("mul.lo.u32 %s, %s, %s;" % (O[j - 1], A[j], bi))
                em.op("mul.hi.u32 %s, %s, %s;" % (O[j], A[j], bi))
        else:
            em.op("add.cc.u32 %s, %s, %s;" % (E[0], E[0], x))   # pending word; its carry enters the odd chain
            for j in range(1, n, 2):
                if j == n - 1:
                    # the top pair of the odd vector is always empty: the carry word of the even chain is NOT folded
                    # back in (a (carry, 0) register pair as 64-bit addend costs ptxas a zeroing move per row and
                    # often splits the wide MAD); it is added once, in the final merge.
                    lo, nw = em.fresh(), em.fresh()
                    em.op("madc.lo.cc.u32 %s, %s, %s, 0;" % (lo, A[j], bi))
                    em.op("madc.hi.u32 %s, %s, %s, 0;" % (nw, A[j], bi))
                    O[j - 1], O[j] = lo, nw
                else:
                    em.op("madc.lo.cc.u32 %s, %s, %s, %s;" % (O[j - 1], A[j], bi, O[j - 1]))
                    em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (O[j], A[j], bi, O[j]))
            for j in range(0, n, 2):
                c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c0, E[j], A[j], bi, E[j]))
                em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], A[j], bi, E[j + 1]))
            tops[i] = em.fresh()                                  # weight 2^(32 (n + i)): limb i of the high half
            em.op("addc.u32 %s, 0, 0;" % tops[i])
        out.append(E[0])                                          # limb i of the product is final
        x = E[1]
        E, O = O, E[2:] + [None, None]
    # remaining n limbs: E + x + O*2^32 + the deferred carry words
    hi = [em.fresh() for _ in range(n)]
    em.op("add.cc.u32 %s, %s, %s;" % (hi[0], E[0], x))
    for k in range(1, n):
        cc = "addc.cc.u32" if k < n - 1 else "addc.u32"
        o = O[k - 1]
        em.op("%s %s, %s, %s;" % (cc, hi[k], E[k], o if o is not None else "0"))
    hi2 = [hi[0]] + [em.fresh() for _ in range(n - 1)]
    for k in range(1, n):
        cc = "add.cc.u32" if k == 1 else ("addc.cc.u32" if k < n - 1 else "addc.u32")
        em.op("%s %s, %s, %s;" % (cc, hi2[k], hi[k], tops[k]))
    return out + hi2


def abs_diff(em, X, Y):
    """(|X - Y| as len(X) names, mask name: 0xffffffff if X < Y else 0)."""
    h = len(X)
    d = [em.fresh() for _ in range(h)]
    e = [em.fresh() for _ in range(h)]
    r = [em.fresh() for _ in range(h)]
    bw = em.fresh()
    for k in range(h):
        em.op("%s %s, %s, %s;" % ("sub.cc.u32" if k == 0 else "subc.cc.u32", d[k], X[k], Y[k]))
    em.op("subc.u32 %s, 0, 0;" % bw)
    for k in range(h):
        cc = "sub.cc.u32" if k == 0 else ("subc.cc.u32" if k < h - 1 else "subc.u32")
        em.op("%s %s, %s, %s;" % (cc, e[k], Y[k], X[k]))
    for k in range(h):
        em.op("lop3.b32 %s, %s, %s, %s, 0xE4;" % (r[k], e[k], d[k], bw))   # bw ? e : d
    return r, bw


def signed_cross(em, zi, zj, mm, sa, sb):
    """zi + zj -+ mm (n limbs each) -> n + 1 limbs: a_i b_j + a_j b_i from the three products of a Karatsuba step;
    mm = |a_i - a_j| |b_i - b_j| is subtracted when the two differences have the same sign (masks sa, sb)."""
    n = len(zi)
    t = [em.fresh() for _ in range(n + 1)]
    for k in range(n):
        em.op("%s %s, %s, %s;" % ("add.cc.u32" if k == 0 else "addc.cc.u32", t[k], zi[k], zj[k]))
    em.op("addc.u32 %s, 0, 0;" % t[n])
    neg = em.fresh()
    em.op("xor.b32 %s, %s, %s;" % (neg, sa, sb))
    em.op("xor.b32 %s, %s, 0xffffffff;" % (neg, neg))
    w = [em.fresh() for _ in range(n)]
    for k in range(n):
        em.op("xor.b32 %s, %s, %s;" % (w[k], mm[k], neg))
    z1 = [em.fresh() for _ in range(n + 1)]
    junk = em.fresh()
    em.op("add.cc.u32 %s, %s, %s;" % (junk, neg, neg))            # carry = 1 iff subtracting (the +1 of -mm)
    for k in range(n):
        em.op("addc.cc.u32 %s, %s, %s;" % (z1[k], t[k], w[k]))
    em.op("addc.u32 %s, %s, %s;" % (z1[n], t[n], neg))
    return z1


def add_at(em, X, Y, off):
    """X + Y * 2^(32 off) for limb-name lists (None = zero limb); the result has max(len) limbs, the final carry is
    dropped (callers know the sum fits)."""
    L = max(len(X), off + len(Y))
    out = []
    started = False
    for k in range(L):
        x = X[k] if k < len(X) else None
        y = Y[k - off] if 0 <= k - off < len(Y) else None
        if not started and y is None:
            out.append(x)
            continue
        r = em.fresh()
        last = k == L - 1
        if not started:
            op = "add.u32" if last else "add.cc.u32"
            started = True
        else:
            op = "addc.u32" if last else "addc.cc.u32"
        em.op("%s %s, %s, %s;" % (op, r, x if x is not None else "0", y if y is not None else "0"))
        out.append(r)
    return out


def three_way_product(em, A, B, sub):
    """Karatsuba with three parts: 6 products of n/3 limbs instead of 9."""
    n = len(A)
    s = n // 3
    a = [A[i * s:(i + 1) * s] for i in range(3)]
    b = [B[i * s:(i + 1) * s] for i in range(3)]
    z = [kara_product(em, a[i], b[i], sub) for i in range(3)]
    cross = {}
    for i, j in ((0, 1), (0, 2), (1, 2)):
        da, sa = abs_diff(em, a[i], a[j])
        db, sb = abs_diff(em, b[i], b[j])
        mm = kara_product(em, da, db, sub)
        cross[(i, j)] = signed_cross(em, z[i], z[j], mm, sa, sb)
    # P = z0 | z1 | z2 (disjoint 2s-limb blocks) + (c01 + c02 B + c12 B^2) B,  B = 2^(32 s)
    M = add_at(em, cross[(0, 1)] + [None] * (2 * s), cross[(0, 2)], s)
    M = add_at(em, M, cross[(1, 2)], 2 * s)
    out = add_at(em, z[0] + z[1] + z[2], M, s)
    assert len(out) == 2 * n
    return out


def kara_product(em, A, B, levels):
    """A*B as 2n limb names. `levels`: 0 = schoolbook rows; k > 0 = k two-way Karatsuba levels; a 3-tuple = one
    two-way level whose three sub-products (low, high, middle) use the given specs; "t3" = one three-way level."""
    n = len(A)
    if levels == "s":
        return school_product(em, A, B)
    if levels == "t3" and n % 6 == 0:
        return three_way_product(em, A, B, 0)
    if isinstance(levels, tuple):
        subs = levels
    elif levels == 0 or n % 4 != 0:
        return school_product(em, A, B)
    else:
        subs = (levels - 1,) * 3
    h = n // 2
    a0, a1, b0, b1 = A[:h], A[h:], B[:h], B[h:]
    da, sa = abs_diff(em, a0, a1)
    db, sb = abs_diff(em, b0, b1)
    z0 = kara_product(em, a0, b0, subs[0])
    z2 = kara_product(em, a1, b1, subs[1])
    mm = kara_product(em, da, db, subs[2])
    # t = z0 + z2 (n limbs + carry limb)
    t = [em.fresh() for _ in range(n + 1)]
    for k in range(n):
        em.op("%s %s, %s, %s;" % ("add.cc.u32" if k == 0 else "addc.cc.u32", t[k], z0[k], z2[k]))
    em.op("addc.u32 %s, 0, 0;" % t[n])
    # z1 = t - mm when the two differences have the same sign, t + mm otherwise (always >= 0, n+1 limbs)
    neg = em.fresh()
    em.op("xor.b32 %s, %s, %s;" % (neg, sa, sb))
    em.op("xor.b32 %s, %s, 0xffffffff;" % (neg, neg))
    w = [em.fresh() for _ in range(n)]
    for k in range(n):
        em.op("xor.b32 %s, %s, %s;" % (w[k], mm[k], neg))
    z1 = [em.fresh() for _ in range(n + 1)]
    junk = em.fresh()
    em.op("add.cc.u32 %s, %s, %s;" % (junk, neg, neg))            # carry = 1 iff subtracting (the +1 of -mm)
    for k in range(n):
        em.op("addc.cc.u32 %s, %s, %s;" % (z1[k], t[k], w[k]))
    em.op("addc.u32 %s, %s, %s;" % (z1[n], t[n], neg))
    # P = z0 + z1 B + z2 B^2
    out = list(z0[:h])
    for k in range(n + 1):
        base = z0[h + k] if k < h else z2[k - h]
        r = em.fresh()
        em.op("%s %s, %s, %s;" % ("add.cc.u32" if k == 0 else "addc.cc.u32", r, base, z1[k]))
        out.append(r)
    for k in range(h + 1, n):
        r = em.fresh()
        cc = "addc.cc.u32" if k < n - 1 else "addc.u32"
        em.op("%s %s, %s, 0;" % (cc, r, z2[k]))
        out.append(r)
    assert len(out) == 2 * n
    return out


def gen_redc(em, N, q, inv, Pw, r_op, inv_opnd=None):
    """r = Pw / R mod p for a 2N-limb Pw < p*R (names), canonical. N*N wide MADs + N MULs. inv_opnd: operand that holds
    inv (default: an immediate)."""
    E = list(Pw[:N])
    O = [None] * N
    x = None
    tops = []
    for i in range(N):
        m = em.fresh()
        if i > 0:
            nw = em.fresh()
            em.op("add.cc.u32 %s, %s, %s;" % (nw, E[0], x))       # carry enters the odd chain
            E[0] = nw
        em.op("mul.lo.u32 %s, %s, %s;" % (m, E[0], inv_opnd or imm(inv)))
        for j in range(1, N, 2):
            lo, hi = em.fresh(), em.fresh()
            if O[j - 1] is None and i == 0:
                em.op("mul.lo.u32 %s, %s, %s;" % (lo, m, imm(q[j])))
                em.op("mul.hi.u32 %s, %s, %s;" % (hi, m, imm(q[j])))
            else:
                c0 = "madc.lo.cc.u32" if (i > 0 or j > 1) else "mad.lo.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c0, lo, m, imm(q[j]), O[j - 1] if O[j - 1] is not None else "0"))
                c1 = "madc.hi.u32" if j == N - 1 else "madc.hi.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c1, hi, m, imm(q[j]), O[j] if O[j] is not None else "0"))
            O[j - 1], O[j] = lo, hi
        nE = list(E)
        for j in range(0, N, 2):
            lo, hi = em.fresh(), em.fresh()
            c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
            em.op("%s %s, %s, %s, %s;" % (c0, lo, m, imm(q[j]), E[j]))
            em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (hi, m, imm(q[j]), E[j + 1]))
            nE[j], nE[j + 1] = lo, hi
        tops.append(em.fresh())                                   # weight 2^(32 (N + i)): limb i after the division by R
        em.op("addc.u32 %s, 0, 0;" % tops[i])
        x = nE[1]
        E, O = O, nE[2:] + [None, None]
    # T = E + x + O*2^32 + deferred carry words <= q, then + P_hi (< q): < 2q
    T = [em.fresh() for _ in range(N)]
    em.op("add.cc.u32 %s, %s, %s;" % (T[0], E[0], x))
    for k in range(1, N):
        cc = "addc.cc.u32" if k < N - 1 else "addc.u32"
        o = O[k - 1]
        em.op("%s %s, %s, %s;" % (cc, T[k], E[k], o if o is not None else "0"))
    V = [em.fresh() for _ in range(N)]
    for k in range(N):
        cc = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < N - 1 else "addc.u32")
        em.op("%s %s, %s, %s;" % (cc, V[k], Pw[N + k], tops[k]))
    U = [em.fresh() for _ in range(N)]
    for k in range(N):
        cc = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < N - 1 else "addc.u32")
        em.op("%s %s, %s, %s;" % (cc, U[k], T[k], V[k]))
    gen_final_sub(em, N, q, U, r_op)


def gen_mul_kara(N, P, inv, a_op, b_op, r_op, levels, inv_opnd=None):
    q = C.limbs32(P, N)
    em = Emit()
    Pw = kara_product(em, [a_op(j) for j in range(N)], [b_op(j) for j in range(N)], levels)
    gen_redc(em, N, q, inv, Pw, r_op, inv_opnd)
    return em


def gen_mulw(N, a_op, b_op, p_op, levels):
    """p (2N limbs) = a*b, unreduced."""
    em = Emit()
    Pw = kara_product(em, [a_op(j) for j in range(N)], [b_op(j) for j in range(N)], levels)
    for k in range(2 * N):
        em.op("mov.u32 %s, %s;" % (p_op(k), Pw[k]))
    return em


def gen_wsubk(N, P, w_op, p_op):
    """w (2N limbs, in/out) = w - p + q^2: the difference of two products of reduced operands, made non-negative by a
    multiple of q; 0 < w < 2 q^2 < q R, so gen_redc's bound holds."""
    K = C.limbs32(P * P, 2 * N)
    em = Emit()
    D = [em.fresh() for _ in range(2 * N)]
    for k in range(2 * N):
        cc = "sub.cc.u32" if k == 0 else ("subc.cc.u32" if k < 2 * N - 1 else "subc.u32")
        em.op("%s %s, %s, %s;" % (cc, D[k], w_op(k), p_op(k)))
    for k in range(2 * N):
        cc = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < 2 * N - 1 else "addc.u32")
        em.op("%s %s, %s, %s;" % (cc, w_op(k), D[k], imm(K[k])))
    return em


def gen_redc_only(N, P, inv, p_op, r_op):
    q = C.limbs32(P, N)
    em = Emit()
    gen_redc(em, N, q, inv, [p_op(k) for k in range(2 * N)], r_op)
    return em


def wrap_io(name, em, outs, ins, inouts, doc):
    """outs / ins / inouts: lists of (c_name, length); asm operands are numbered outs, inouts, ins in that order."""
    o = ["// %s" % doc]
    args = ["uint32_t (&%s)[%d]" % (c, n) for c, n in outs + inouts] + ["const uint32_t (&%s)[%d]" % (c, n) for c, n in ins]
    o.append("__device__ __forceinline__ void %s(%s)" % (name, ", ".join(args)))
    o.append("{")
    o.append("    asm volatile(")
    o.append('        "{\\n\\t"')
    o.append('        ".reg .u32 v<%d>;\\n\\t"' % max(em.nreg, 1))
    o.append('        ".reg .pred pb;\\n\\t"')
    for ln in em.lines:
        o.append('        "%s\\n\\t"' % ln)
    o.append('        "}\\n\\t"')
    cons = ['"=r"(%s[%d])' % (c, k) for c, n in outs for k in range(n)] + ['"+r"(%s[%d])' % (c, k) for c, n in inouts for k in range(n)]
    o.append("        : %s" % ", ".join(cons))
    o.append("        : %s);" % ", ".join('"r"(%s[%d])' % (c, k) for c, n in ins for k in range(n)))
    o.append("}")
    o.append("")
    return "\n".join(o)


# Karatsuba levels of <tag>_mul_regs (0 = the interleaved CIOS form). Override: ZECALE_FQ_KARA / ZECALE_FR_KARA.
def _kara_spec(v):
    """"1" -> 1, "t3" -> "t3", "1,0,0" -> (1, 0, 0), "s" -> separated schoolbook product + reduction (no Karatsuba)."""
    if v in ("t3", "s"):
        return v
    if "," in v:
        return tuple(int(x) for x in v.split(","))
    return int(v)


KARA_LEVELS = {"fq": _kara_spec(os.environ.get("ZECALE_FQ_KARA", "1")), "fr": _kara_spec(os.environ.get("ZECALE_FR_KARA", "0"))}


Q_FROM_REGS = {"fr": int(os.environ.get("ZECALE_FR_QREGS", "2"))}      # 1: modulus limbs and inv from registers, 2: inv only


def n_wide(em):
    """wide MADs of a body (lo/hi pairs fuse into one IMAD.WIDE) + the single-word MULs by inv."""
    return sum(1 for l in em.lines if l.startswith(("mul.lo", "mad.lo", "madc.lo")))


def wrap(name, N, em, n_in, doc):
    """Emit a __forceinline__ device function around one asm block."""
    out = []
    out.append("// %s" % doc)
    args = ["uint32_t (&r)[%d]" % N] + ["const uint32_t (&%s)[%d]" % (c, N) for c in "ab"[:n_in]]
    out.append("__device__ __forceinline__ void %s(%s)" % (name, ", ".join(args)))
    out.append("{")
    out.append("    asm volatile(")
    out.append('        "{\\n\\t"')
    out.append('        ".reg .u32 v<%d>;\\n\\t"' % em.nreg)
    out.append('        ".reg .pred pb;\\n\\t"')
    for ln in em.lines:
        out.append('        "%s\\n\\t"' % ln)
    out.append('        "}\\n\\t"')
    outs = ", ".join('"=r"(r[%d])' % k for k in range(N))
    ins = ", ".join('"r"(%s[%d])' % (c, k) for c in "ab"[:n_in] for k in range(N))
    out.append("        : %s" % outs)
    out.append("        : %s);" % ins)
    out.append("}")
    out.append("")
    return "\n".join(out)


def gen_field(tag, N, P, inv):
    r_op = lambda k: "%%%d" % k
    a_op = lambda k: "%%%d" % (N + k)
    b_op = lambda k: "%%%d" % (2 * N + k)
    s = []
    lv = KARA_LEVELS.get(tag, 0)
    if lv and Q_FROM_REGS.get(tag):
        # Karatsuba product + separate reduction with inv as an operand (see the Q_FROM_REGS branch below)
        em = gen_mul_kara(N, P, inv, a_op, b_op, r_op, lv, inv_opnd="%%%d" % (3 * N))
        s.append(wrap_io("%s_mul_q_regs" % tag, em, [("r", N)], [("a", N), ("b", N), ("q", 1)], [],
                         "r = a*b/R mod p: Karatsuba product (spec %s) + word-serial reduction, %d wide MADs; q = {inv}." % (lv, n_wide(em) - N)))
        s.append("__device__ __forceinline__ void %s_mul_regs(uint32_t (&r)[%d], const uint32_t (&a)[%d], const uint32_t (&b)[%d])\n"
                 "{\n    uint32_t q[1];\n    q[0] = %s_INV32_DEV;\n    %s_mul_q_regs(r, a, b, q);\n}\n" % (tag, N, N, N, tag.upper(), tag))
    elif lv:
        em = gen_mul_kara(N, P, inv, a_op, b_op, r_op, lv)
        s.append(wrap("%s_mul_regs" % tag, N, em, 2,
                      "r = a*b/R mod p  (Montgomery product, fully reduced): Karatsuba product (spec %s) + word-serial "
                      "reduction, %d wide MADs." % (lv, n_wide(em) - N)))
        pc = lambda base: (lambda k: "%%%d" % (base + k))
        s.append(wrap_io("%s_mulw_regs" % tag, gen_mulw(N, pc(2 * N), pc(3 * N), pc(0), lv), [("p", 2 * N)], [("a", N), ("b", N)], [],
                         "p = a*b, 2N limbs, unreduced (the product half of %s_mul_regs)." % tag))
        s.append(wrap_io("%s_wsubk_regs" % tag, gen_wsubk(N, P, pc(0), pc(2 * N)), [], [("p", 2 * N)], [("w", 2 * N)],
                         "w = w - p + q^2 (difference of two unreduced products, kept in [0, qR) for %s_redc_regs)." % tag))
        s.append(wrap_io("%s_redc_regs" % tag, gen_redc_only(N, P, inv, pc(N), pc(0)), [("r", N)], [("p", 2 * N)], [],
                         "r = p/R mod q for p < qR, canonical (the reduction half of %s_mul_regs)." % tag))
        s.append(wrap("%s_mul_cios_regs" % tag, N, gen_mul(N, P, inv, a_op, b_op, r_op), 2,
                      "the interleaved (CIOS) form of the same product, %d wide MADs; kept for the field tests." % (2 * N * N)))
    elif Q_FROM_REGS.get(tag):
        # inv (mode 2), or inv and the modulus limbs (mode 1), reach the body through registers / the constant bank instead of
        # as immediates. Fr has inv = 0xffffffff: told so, ptxas replaces m = T0 * inv by a negation and then splits every wide MAD
        # of the m*q chains into IMAD.X + IMAD.HI.U32.X (6 pipe cycles for 4; k_ntt_pass: 1815 -> 1348 FMA-pipe instructions).
        inv_only = Q_FROM_REGS[tag] == 2
        nq = 1 if inv_only else N + 1
        q_op = (lambda k: "%%%d" % (3 * N)) if inv_only else (lambda k: "%%%d" % (3 * N + k))
        em = gen_mul(N, P, inv, a_op, b_op, r_op, q_op=q_op, inv_only=inv_only)
        s.append(wrap_io("%s_mul_q_regs" % tag, em, [("r", N)], [("a", N), ("b", N), ("q", nq)], [],
                         "r = a*b/R mod p (Montgomery product, fully reduced), %d wide MADs; q = {inv} or {modulus limbs, inv}, "
                         "opaque to the compiler." % (2 * N * N)))
        fill = "" if inv_only else "#pragma unroll\n    for (int k = 0; k < %d; k++) q[k] = %s_P[k];\n" % (N, tag.upper())
        s.append("__device__ __forceinline__ void %s_mul_regs(uint32_t (&r)[%d], const uint32_t (&a)[%d], const uint32_t (&b)[%d])\n"
                 "{\n    uint32_t q[%d];\n%s    q[%d] = %s_INV32_DEV;\n    %s_mul_q_regs(r, a, b, q);\n}\n"
                 % (tag, N, N, N, nq, fill, nq - 1, tag.upper(), tag))
    else:
        s.append(wrap("%s_mul_regs" % tag, N, gen_mul(N, P, inv, a_op, b_op, r_op), 2,
                      "r = a*b/R mod p  (Montgomery product, fully reduced). %d wide MADs." % (2 * N * N)))
    s.append(wrap("%s_sqr_regs" % tag, N, gen_mul(N, P, inv, a_op, a_op, r_op, square=True), 1,
                  "r = a*a/R mod p. %d wide MADs (even columns below the diagonal dropped)." % (2 * N * N - N * N // 4)))
    s.append(wrap("%s_add_regs" % tag, N, gen_add(N, P, a_op, b_op, r_op), 2, "r = a+b mod p."))
    s.append(wrap("%s_sub_regs" % tag, N, gen_sub(N, P, a_op, b_op, r_op), 2, "r = a-b mod p."))
    return "\n".join(s)


HEADER = """// GENERATED by tools/gen_field_asm.py — do not edit by hand.
// BW6-761 Fq (24 x u32) and Fr (12 x u32) Montgomery arithmetic as PTX carry chains.
// See the generator's docstring and DESIGN.md §3.1 for the derivation and bounds.
#pragma once
#include <cstdint>

#include "field_consts.cuh"

namespace zecale {

"""


def carr(name, vals):
    body = ", ".join("0x%08xu" % v for v in vals)
    return "__device__ __constant__ uint32_t %s[%d] = {%s};\n" % (name, len(vals), body)


def gen_consts():
    """Constants header: moduli, Montgomery one / R^2, p-2 exponent bits, Fr NTT roots."""
    o = "// GENERATED by tools/gen_field_asm.py — do not edit by hand.\n"
    o += "// Field constants derived in tools/bw6_constants.py from the two moduli of\n"
    o += "// libff bw6_761_init.cpp:36-146 (cross-checked there against the reference's inv/R^2).\n"
    o += "#pragma once\n#include <cstdint>\nnamespace zecale {\n"
    for tag, P, N, K in (("FQ", C.Q, C.FQ_LIMBS32, C.FQ), ("FR", C.R, C.FR_LIMBS32, C.FR)):
        o += carr("%s_P" % tag, C.limbs32(P, N))
        o += carr("%s_ONE" % tag, C.limbs32(K["one"], N))
        # host-side copy (the runtime fills zero points / unit coset factors without a device round trip)
        o += "static const uint32_t %s_ONE_HOST[%d] = {%s};\n" % (tag, N, ", ".join("0x%08xu" % v for v in C.limbs32(K["one"], N)))
        o += carr("%s_R2" % tag, C.limbs32(K["r2"], N))
        o += carr("%s_PM2" % tag, C.limbs32(P - 2, N))
        o += "static constexpr uint32_t %s_INV32 = 0x%08xu;\n" % (tag, K["inv32"])
        o += "__device__ __constant__ uint32_t %s_INV32_DEV = 0x%08xu;\n" % (tag, K["inv32"])
        o += "static constexpr int %s_BITS = %d;\n" % (tag, P.bit_length())
    # 2-adic root of unity of order 2^46 and its inverse, Montgomery form
    Rm = (1 << C.FR_RBITS) % C.R
    w = C.FR_ROOT_OF_UNITY
    o += carr("FR_ROOT_2S", C.limbs32(w * Rm % C.R, 12))
    o += carr("FR_ROOT_2S_INV", C.limbs32(pow(w, -1, C.R) * Rm % C.R, 12))
    o += "static constexpr int FR_TWO_ADICITY = %d;\n" % C.FR_S
    Rq = (1 << C.FQ_RBITS) % C.Q
    o += carr("FQ_R3", C.limbs32(C.FQ["r3"], 24))
    for nm, pt in (("G1_GEN", C.G1_GEN), ("G2_GEN", C.G2_GEN)):
        o += carr(nm + "_X", C.limbs32(pt[0] * Rq % C.Q, 24))
        o += carr(nm + "_Y", C.limbs32(pt[1] * Rq % C.Q, 24))
    o += "} // namespace zecale\n"
    return o


def main():
    dstc = os.path.join(os.path.dirname(__file__), "..", "zecale_b200", "csrc", "field_consts.cuh")
    if not os.environ.get("ZECALE_FIELD_ASM_OUT"):
        with open(dstc, "w") as f:
            f.write(gen_consts())
    out = HEADER
    out += gen_field("fq", C.FQ_LIMBS32, C.Q, C.FQ["inv32"])
    out += gen_field("fr", C.FR_LIMBS32, C.R, C.FR["inv32"])
    out += "\n} // namespace zecale\n"
    dst = os.environ.get("ZECALE_FIELD_ASM_OUT") or os.path.join(os.path.dirname(__file__), "..", "zecale_b200", "csrc", "field_asm.cuh")
    with open(dst, "w") as f:
        f.write(out)
    print("wrote", os.path.abspath(dst), len(out.splitlines()), "lines")


if __name__ == "__main__":
    main()
