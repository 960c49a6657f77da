"""Generator, integer model and checker of the tensor-core Montgomery product prototype (scripts/probes/imma_mul.cu).

Idea. A Montgomery product r = a b R^-1 mod q is   P = a b;  m = (P mod R) q' mod R;  r = (P + m q) / R  (q' = -1/q mod R).
Only P multiplies two variables. The other two products multiply by CONSTANTS, i.e. they are matrix products of the
operand's byte vector with a constant Toeplitz matrix: column c of (x * k) = sum_i x.byte[i] * k.byte[c - i]. For the 32
multiplications a warp holds (one per lane) that is a [32 x 96] x [96 x 96] and a [32 x 96] x [96 x 192] u8 GEMM with
s32 accumulation (96 * 255^2 < 2^23): 144 mma.sync.m16n8k32.u8 per warp in place of 600 of the 1176 IMAD.WIDE per lane.

This file writes imma_consts.h (the B-operand fragments of the two Toeplitz matrices, one 256-byte tile per diagonal
offset, and the PTX text of the 24x24-limb product a*b as even/odd carry chains without reduction) and checks dumps."""
import os
import struct
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "tools"))
import bw6_constants as C
import ptx_sim

N = 24
Q = C.Q
R = 1 << (32 * N)
QP = (-pow(Q, -1, R)) % R          # q' = -1/q mod R


def byte(x, i):
    return (x >> (8 * i)) & 0xFF if 0 <= i < 96 else 0


def b_tile(const, d):
    """B fragment (b0, b1) per lane of the 32 x 8 tile whose element (k, n) is const.byte[8 d + n - k]
    (mma.m16n8k32 .col B: lane l holds rows 4 (l % 4) .. + 3 and 16 + 4 (l % 4) .. + 3 of column l / 4)."""
    out = []
    for l in range(32):
        n, k0 = l // 4, 4 * (l % 4)
        regs = []
        for base in (0, 16):
            v = 0
            for t in range(4):
                v |= byte(const, 8 * d + n - (base + k0 + t)) << (8 * t)
            regs.append(v)
        out.append(regs)
    return out


# ---- a * b (48 limbs) as even / odd carry chains: gen_field_asm.py's row structure without the m*q half
def gen_product(a_op, b_op, p_op, zero_op=None, depth=1):
    """zero_op: an operand that is zero at run time but opaque to ptxas. Row i + depth multiplies by
    b ^ (top_i & zero), so it cannot start before row i has finished: without the Montgomery reduction's serial
    dependency ptxas pipelines many rows at once, runs out of carry predicates and spills them bit by bit."""
    lines, cnt = [], [0]
    tops = {}

    def fresh():
        cnt[0] += 1
        return "w%d" % (cnt[0] - 1)

    E = [fresh() for _ in range(N)]
    O = [fresh() for _ in range(N)]
    x = None
    for i in range(N):
        bi = b_op(i)
        if zero_op is not None and i - depth in tops:
            t, bi = fresh(), fresh()
            lines.append("and.b32 %s, %s, %s;" % (t, tops[i - depth], zero_op))
            lines.append("xor.b32 %s, %s, %s;" % (bi, b_op(i), t))
        if i == 0:
            for j in range(0, N, 2):
                lines.append("mul.lo.u32 %s, %s, %s;" % (E[j], a_op(j), bi))
                lines.append("mul.hi.u32 %s, %s, %s;" % (E[j + 1], a_op(j), bi))
            for j in range(1, N, 2):
                lines.append("mul.lo.u32 %s, %s, %s;" % (O[j - 1], a_op(j), bi))
                lines.append("mul.hi.u32 %s, %s, %s;" % (O[j], a_op(j), bi))
            top = fresh()
            lines.append("add.u32 %s, 0, 0;" % top)
        else:
            lines.append("add.cc.u32 %s, %s, %s;" % (E[0], E[0], x))   # fold the pending word; carry enters the odd chain
            for j in range(1, N, 2):
                lines.append("madc.lo.cc.u32 %s, %s, %s, %s;" % (O[j - 1], a_op(j), bi, O[j - 1]))
                if j == N - 1:
                    nw = fresh()
                    lines.append("madc.hi.u32 %s, %s, %s, 0;" % (nw, a_op(j), bi))
                    O[j] = nw
                else:
                    lines.append("madc.hi.cc.u32 %s, %s, %s, %s;" % (O[j], a_op(j), bi, O[j]))
            for j in range(0, N, 2):
                c0 = "mad.lo.cc.u32" if j == 0 else "madc.lo.cc.u32"
                lines.append("%s %s, %s, %s, %s;" % (c0, E[j], a_op(j), bi, E[j]))
                lines.append("madc.hi.cc.u32 %s, %s, %s, %s;" % (E[j + 1], a_op(j), bi, E[j + 1]))
            top = fresh()
            lines.append("addc.u32 %s, 0, 0;" % top)
        tops[i] = top
        lines.append("add.u32 %s, %s, 0;" % (p_op(i), E[0]))           # limb i of the product is final
        x = E[1]
        E, O = O, E[2:] + [top, None]
    lines.append("add.cc.u32 %s, %s, %s;" % (p_op(N), E[0], x))
    for k in range(1, N):
        cc = "addc.cc.u32" if k < N - 1 else "addc.u32"
        lines.append("%s %s, %s, %s;" % (cc, p_op(N + k), E[k], O[k - 1]))
    return lines, cnt[0]


def gen_normalise(L, limb_op, dig_op, cin_op, cout_op):
    """limbs[L] = sum of 2 L radix-2^16 digits (each < 2^32) + cin, carry out in cout. The digit pairs are folded in
    parallel (lo_k = d_2k + (d_2k+1 << 16), hi_k = (d_2k+1 >> 16) + carry); only one add-with-carry per limb is serial."""
    lines = []
    for k in range(L):
        lines.append("shl.b32 t%d, %s, 16;" % (k, dig_op(2 * k + 1)))
        lines.append("add.cc.u32 l%d, %s, t%d;" % (k, dig_op(2 * k), k))
        lines.append("shr.b32 h%d, %s, 16;" % (k, dig_op(2 * k + 1)))
        lines.append("addc.u32 h%d, h%d, 0;" % (k, k))
    lines.append("add.cc.u32 %s, l0, %s;" % (limb_op(0), cin_op))
    for k in range(1, L):
        lines.append("addc.cc.u32 %s, l%d, h%d;" % (limb_op(k), k, k - 1))
    lines.append("addc.u32 %s, h%d, 0;" % (cout_op, L - 1))
    return lines


def check_normalise_ptx():
    import random
    rng = random.Random(9)
    L = 24
    lines = gen_normalise(L, lambda k: "r%d" % k, lambda i: "d%d" % i, "cin", "cout")
    for _ in range(50):
        regs = {"d%d" % i: rng.choice((rng.randrange(1 << 32), 0xFFFFFFFF, 0)) for i in range(2 * L)}
        regs["cin"] = rng.randrange(1 << 18)
        want = sum(regs["d%d" % i] << (16 * i) for i in range(2 * L)) + regs["cin"]
        ptx_sim.run(lines, regs)
        got = sum(regs["r%d" % k] << (32 * k) for k in range(L)) + (regs["cout"] << (32 * L))
        assert got == want
    print("normalise PTX ok (%d lines)" % len(lines))


def check_product_ptx():
    import random
    rng = random.Random(3)
    lines, _ = gen_product(lambda j: "a%d" % j, lambda i: "b%d" % i, lambda k: "p%d" % k, "0x0")
    for a, b in [(R - 1, R - 1), (0, 5), (1, R - 1)] + [(rng.randrange(R), rng.randrange(R)) for _ in range(50)]:
        regs = {}
        for k in range(N):
            regs["a%d" % k] = (a >> (32 * k)) & 0xFFFFFFFF
            regs["b%d" % k] = (b >> (32 * k)) & 0xFFFFFFFF
        ptx_sim.run(lines, regs)
        got = sum(regs["p%d" % k] << (32 * k) for k in range(2 * N))
        assert got == a * b, (hex(a), hex(b))
    print("product PTX ok (%d lines)" % len(lines))


# ---- the kernel's data flow, lane by lane, on Python integers
def model_mul(xs, ys):
    """32 multiplications at once, through the same fragments, column pairs and digit normalisation as the kernel."""
    tq, tn = [b_tile(QP, d) for d in range(16)], [b_tile(Q, d) for d in range(16)]

    def gemm(rows, tiles, n_tiles):
        """rows: 32 ints < R. Returns per row the list of 4 * n_tiles radix-2^16 digits (each < 2^32)."""
        dig = [[0] * (4 * n_tiles) for _ in range(32)]
        for nt in range(n_tiles):
            for mt in range(2):
                acc = [[0, 0, 0, 0] for _ in range(32)]
                for ks in range(3):
                    d = nt - 4 * ks
                    if not 0 <= d < 16:
                        continue
                    for l in range(32):
                        g, qd = l // 4, l % 4
                        # D[row][n] += sum_k A[row][k] B[k][n]; lane l owns D rows g, g + 8 and columns 2 qd, 2 qd + 1
                        for ci, (row, col) in enumerate(((g, 2 * qd), (g, 2 * qd + 1), (g + 8, 2 * qd), (g + 8, 2 * qd + 1))):
                            v = rows[16 * mt + row]
                            s = 0
                            for k in range(32):
                                # B[k][col] as stored in the fragment table of lane 4 * col + (k % 16) / 4
                                src = tiles[d][4 * col + (k % 16) // 4][k // 16]
                                s += byte(v, 32 * ks + k) * ((src >> (8 * (k % 4))) & 0xFF)
                            acc[l][ci] += s
                for l in range(32):
                    g, qd = l // 4, l % 4
                    assert max(acc[l]) < 1 << 23
                    dig[16 * mt + g][4 * nt + qd] = acc[l][0] + (acc[l][1] << 8)
                    dig[16 * mt + g + 8][4 * nt + qd] = acc[l][2] + (acc[l][3] << 8)
        return dig

    def normalise(digs):
        limbs, carry = [], 0
        for k in range(len(digs) // 2):
            t = digs[2 * k] + (digs[2 * k + 1] << 16) + carry
            limbs.append(t & 0xFFFFFFFF)
            carry = t >> 32
        return limbs, carry

    P = [x * y for x, y in zip(xs, ys)]
    m = []
    for digs in gemm([p % R for p in P], tq, 12):
        limbs, _ = normalise(digs)
        m.append(sum(v << (32 * k) for k, v in enumerate(limbs)))
    out = []
    for p, mm, digs in zip(P, m, gemm(m, tn, 24)):
        assert mm == (p % R) * QP % R
        limbs, carry = normalise(digs)
        assert carry == 0
        U = sum(v << (32 * k) for k, v in enumerate(limbs))
        assert U == mm * Q
        r = (p >> (32 * N)) + (U >> (32 * N)) + (1 if p % R else 0)
        assert r == (p + U) // R
        out.append(r - Q if r >= Q else r)
    return out


def selftest():
    import random
    rng = random.Random(5)
    check_product_ptx()
    check_normalise_ptx()
    xs = [rng.randrange(Q) for _ in range(32)]
    ys = [rng.randrange(Q) for _ in range(32)]
    xs[0], ys[1], xs[2], ys[2] = 0, 0, Q - 1, Q - 1
    rinv = pow(R, -1, Q)
    for x, y, r in zip(xs, ys, model_mul(xs, ys)):
        assert r == x * y * rinv % Q
    print("model ok")


def header():
    o = ["// generated by scripts/probes/imma_model.py", "#pragma once", "#include <cstdint>"]
    for name, const in (("IMMA_TQ", QP), ("IMMA_TN", Q)):
        rows = []
        for d in range(16):
            t = b_tile(const, d)
            rows.append("{" + ", ".join("{0x%08xu, 0x%08xu}" % (r[0], r[1]) for r in t) + "}")
        o.append("static const uint32_t %s_HOST[16][32][2] = {\n%s};" % (name, ",\n".join(rows)))
    o.append("static const uint32_t IMMA_Q_HOST[24] = {%s};" % ", ".join("0x%08xu" % v for v in C.limbs32(Q, N)))
    o.append("static const uint32_t IMMA_ONE_HOST[24] = {%s};" % ", ".join("0x%08xu" % v for v in C.limbs32(R % Q, N)))
    for depth in (1, 2, 3):
        lines, nreg = gen_product(lambda j: "%%%d" % (2 * N + j), lambda i: "%%%d" % (3 * N + i), lambda k: "%%%d" % k, "%%%d" % (4 * N), depth)
        o.append("// p[48] = a[24] * b[24]: 576 wide MADs, even / odd carry chains, no reduction; at most %d rows in flight" % depth)
        o.append("// zero: 0 at run time, unknown to the compiler (serialises the rows, see gen_product)")
        o.append("__device__ __forceinline__ void imma_full_product_d%d(uint32_t (&p)[48], const uint32_t (&a)[24], const uint32_t (&b)[24], uint32_t zero)" % depth)
        o.append("{")
        o.append("    asm volatile(")
        o.append('        "{\\n\\t"')
        o.append('        ".reg .u32 w<%d>;\\n\\t"' % nreg)
        for ln in lines:
            o.append('        "%s\\n\\t"' % ln)
        o.append('        "}\\n\\t"')
        o.append("        : " + ", ".join('"=&r"(p[%d])' % k for k in range(2 * N)))
        o.append("        : " + ", ".join('"r"(%s[%d])' % (c, k) for c in "ab" for k in range(N)) + ', "r"(zero));')
        o.append("}")
    L = 24
    lines = gen_normalise(L, lambda k: "%%%d" % k, lambda i: "%%%d" % (L + 1 + i), "%%%d" % (3 * L + 1), "%%%d" % L)
    o.append("// limbs[24] = 48 radix-2^16 digits (each < 2^32) + cin; returns the carry out. One serial add-with-carry per limb.")
    o.append("__device__ __forceinline__ uint32_t imma_normalise24(uint32_t (&limbs)[24], const uint32_t (&d)[48], uint32_t cin)")
    o.append("{")
    o.append("    uint32_t cout;")
    o.append("    asm volatile(")
    o.append('        "{\\n\\t"')
    o.append('        ".reg .u32 t<%d>, l<%d>, h<%d>;\\n\\t"' % (L, L, L))
    for ln in lines:
        o.append('        "%s\\n\\t"' % ln)
    o.append('        "}\\n\\t"')
    o.append("        : " + ", ".join('"=&r"(limbs[%d])' % k for k in range(L)) + ', "=&r"(cout)')
    o.append("        : " + ", ".join('"r"(d[%d])' % i for i in range(2 * L)) + ', "r"(cin));')
    o.append("    return cout;")
    o.append("}")
    return "\n".join(o) + "\n"


def check(path):
    """Dump written by imma_mul: sections of (n, iters) then n * (a[24], b[24], out[24]) u32 limbs."""
    raw = open(path, "rb").read()
    pos, bad_total = 0, 0
    rinv = pow(R, -1, Q)
    while pos < len(raw):
        n, iters = struct.unpack_from("<2Q", raw, pos)
        pos += 16
        bad = 0
        for _ in range(n):
            v = struct.unpack_from("<72I", raw, pos)
            pos += 72 * 4
            a, b, o = (sum(x << (32 * k) for k, x in enumerate(v[s:s + 24])) for s in (0, 24, 48))
            want = a
            for _ in range(iters):
                want = want * b * rinv % Q
            if o != want % Q:
                bad += 1
        print('{"probe": "imma_mul exactness vs python integers (canonical result < q)", "chained_products": %d, "cases": %d, '
              '"mismatches": %d}' % (iters, n, bad))
        bad_total += bad
    return bad_total


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "header":
        open(os.path.join(HERE, "imma_consts.h"), "w").write(header())
    elif len(sys.argv) > 2 and sys.argv[1] == "check":
        sys.exit(1 if check(sys.argv[2]) else 0)
    else:
        selftest()
