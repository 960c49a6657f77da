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


def gen_mul(N, P, inv, a_op, b_op, r_op, square=False):
    """a_op(j), b_op(i), r_op(k) -> PTX operand strings (%k placeholders)."""
    q = C.limbs32(P, N)
    em = Emit()
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
            em.op("mul.lo.u32 %s, %s, %s;" % (m, E[0], imm(inv)))
            # even q chain
            for j in range(0, N, 2):
                c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c0, E[j], m, imm(q[j]), E[j]))
                em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], m, imm(q[j]), E[j + 1]))
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
            em.op("mul.lo.u32 %s, %s, %s;" % (m, E[0], imm(inv)))
            for j in range(0, N, 2):
                c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
                em.op("%s %s, %s, %s, %s;" % (c0, E[j], m, imm(q[j]), E[j]))
                em.op("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], m, imm(q[j]), E[j + 1]))
            em.op("addc.u32 %s, %s, 0;" % (top, top))
        # odd q chain (never carries out)
        for j in range(1, N, 2):
            c0 = "mad.lo.cc.u32" if j == 1 else "madc.lo.cc.u32"
            em.op("%s %s, %s, %s, %s;" % (c0, O[j - 1], m, imm(q[j]), O[j - 1]))
            c1 = "madc.hi.u32" if j == N - 1 else "madc.hi.cc.u32"
            em.op("%s %s, %s, %s, %s;" % (c1, O[j], m, imm(q[j]), O[j]))
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
    with open(dstc, "w") as f:
        f.write(gen_consts())
    out = HEADER
    out += gen_field("fq", C.FQ_LIMBS32, C.Q, C.FQ["inv32"])
    out += gen_field("fr", C.FR_LIMBS32, C.R, C.FR["inv32"])
    out += "\n} // namespace zecale\n"
    dst = os.path.join(os.path.dirname(__file__), "..", "zecale_b200", "csrc", "field_asm.cuh")
    with open(dst, "w") as f:
        f.write(out)
    print("wrote", os.path.abspath(dst), len(out.splitlines()), "lines")


if __name__ == "__main__":
    main()
