// BW6-761 G1 / G2 point arithmetic for sm_100a.
//
// Both groups are short-Weierstrass curves y^2 = x^3 + b over the SAME base field Fq
// with a = 0 (G1: b = -1, G2: M-twist b' = 4; libff bw6_761_init.cpp:264-281), and b
// never enters an addition or doubling formula, so one set of device functions serves
// both groups (cf. libff bw6_761_g1.cpp:111-358 and bw6_761_g2.cpp:116-364, which are
// the same formulas twice).
//
// Device representations
//   Affine   {x, y}            192 B, the proving-key image; (0,0) encodes infinity
//                              (never a curve point for b != 0). Built from libff's
//                              "special" form {X, Y, Z in {0, one}} by ec_import_special.
//   XYZZ     {x, y, zz, zzz}   384 B accumulators, x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2,
//                              infinity <=> zz == 0.
// libff keeps homogeneous projective {X,Y,Z} (bw6_761_g1.hpp:14-91). The coordinate
// system is free: parity is on the normalised affine result, which is unique.
//
// Costs (Fq products, S counted as M): madd 8M+2S (libff mixed_add: 9M+2S),
// add 12M+2S (libff operator+: 4M pre-check + 12M+2S), dbl 6M+3S (libff dbl: 5M+6S).
#pragma once
#include "field.cuh"

namespace zecale {

struct alignas(16) Affine {
    Fq x, y;
};
struct alignas(16) XYZZ {
    Fq x, y, zz, zzz;
};
// libff in-memory point: homogeneous projective {X, Y, Z}
struct alignas(8) Proj {
    Fq X, Y, Z;
};
static_assert(sizeof(Affine) == 192 && sizeof(XYZZ) == 384 && sizeof(Proj) == 288, "point layout");

__device__ __forceinline__ bool affine_is_inf(const Affine &p) { return fq_is_zero(p.x) && fq_is_zero(p.y); }
__device__ __forceinline__ bool xyzz_is_inf(const XYZZ &p) { return fq_is_zero(p.zz); }

__device__ __forceinline__ void xyzz_set_inf(XYZZ &p)
{
    fq_set_zero(p.x);
    fq_set_one(p.y);
    fq_set_zero(p.zz);
    fq_set_zero(p.zzz);
}

__device__ __forceinline__ void xyzz_copy(XYZZ &r, const XYZZ &a)
{
    fq_copy(r.x, a.x);
    fq_copy(r.y, a.y);
    fq_copy(r.zz, a.zz);
    fq_copy(r.zzz, a.zzz);
}

// acc = 2 * (px, py) for an affine point (mdbl-2008-s-1, a = 0).
__device__ __noinline__ void xyzz_mdbl(XYZZ &acc, const Fq &px, const Fq &py)
{
    Fq u, v, w, s, m, t;
    fq_dbl(u, py);          // U = 2*Y1
    fq_sqr(v, u);           // V = U^2
    fq_mul(w, u, v);        // W = U*V
    fq_mul(s, px, v);       // S = X1*V
    fq_sqr(t, px);
    fq_dbl(m, t);
    fq_add(m, m, t);        // M = 3*X1^2
    fq_sqr(t, m);
    fq_sub(t, t, s);
    fq_sub(acc.x, t, s);    // X3 = M^2 - 2S
    fq_sub(t, s, acc.x);
    fq_mul(t, m, t);
    fq_mul(u, w, py);
    fq_sub(acc.y, t, u);    // Y3 = M*(S-X3) - W*Y1
    fq_copy(acc.zz, v);
    fq_copy(acc.zzz, w);
}

// acc = 2*acc (dbl-2008-s-1, a = 0). 6M + 3S.
__device__ __noinline__ void xyzz_dbl(XYZZ &acc)
{
    if (xyzz_is_inf(acc)) return;
    Fq u, v, w, s, m, t;
    fq_dbl(u, acc.y);
    fq_sqr(v, u);
    fq_mul(w, u, v);
    fq_mul(s, acc.x, v);
    fq_sqr(t, acc.x);
    fq_dbl(m, t);
    fq_add(m, m, t);
    fq_sqr(t, m);
    fq_sub(t, t, s);
    fq_sub(acc.x, t, s);     // X3
    fq_sub(t, s, acc.x);
    fq_mul(t, m, t);
    fq_mul(u, w, acc.y);
    fq_sub(acc.y, t, u);     // Y3
    fq_mul(acc.zz, v, acc.zz);
    fq_mul(acc.zzz, w, acc.zzz);
}

// acc += (px, ±py)  — mixed addition, madd-2008-s, 8M + 2S.
// Handles exactly the cases libff's mixed_add does (bw6_761_g1.cpp:263-316):
// either operand at infinity, P == Q (doubling), P == -Q (infinity).
__device__ __noinline__ void xyzz_madd(XYZZ &acc, const Affine &p, bool negate)
{
    if (affine_is_inf(p)) return;
    Fq py;
    if (negate)
        fq_neg(py, p.y);
    else
        fq_copy(py, p.y);
    if (xyzz_is_inf(acc)) {
        fq_copy(acc.x, p.x);
        fq_copy(acc.y, py);
        fq_set_one(acc.zz);
        fq_set_one(acc.zzz);
        return;
    }
    Fq pp, r, t, q;
    fq_mul(pp, p.x, acc.zz);     // U2
    fq_mul(r, py, acc.zzz);      // S2
    fq_sub(pp, pp, acc.x);       // P = U2 - X1
    fq_sub(r, r, acc.y);         // R = S2 - Y1
    if (fq_is_zero(pp)) {
        if (fq_is_zero(r))
            xyzz_mdbl(acc, p.x, py);
        else
            xyzz_set_inf(acc);
        return;
    }
    fq_sqr(t, pp);               // PP
    fq_mul(pp, pp, t);           // PPP
    fq_mul(q, acc.x, t);         // Q = X1*PP
    fq_mul(acc.zz, acc.zz, t);   // ZZ3 = ZZ1*PP
    fq_mul(acc.zzz, acc.zzz, pp);// ZZZ3 = ZZZ1*PPP
    fq_sqr(t, r);
    fq_sub(t, t, pp);
    fq_sub(t, t, q);
    fq_sub(acc.x, t, q);         // X3 = R^2 - PPP - 2Q
    fq_sub(q, q, acc.x);
    fq_mul(q, r, q);
    fq_mul(t, acc.y, pp);
    fq_sub(acc.y, q, t);         // Y3 = R*(Q-X3) - Y1*PPP
}

// acc += b — general addition, add-2008-s, 12M + 2S, with the same special cases
// as libff's operator+ (bw6_761_g1.cpp:111-194).
__device__ __noinline__ void xyzz_add(XYZZ &acc, const XYZZ &b)
{
    if (xyzz_is_inf(b)) return;
    if (xyzz_is_inf(acc)) {
        xyzz_copy(acc, b);
        return;
    }
    Fq u1, s1, pp, r, t;
    fq_mul(u1, acc.x, b.zz);     // U1
    fq_mul(pp, b.x, acc.zz);     // U2
    fq_mul(s1, acc.y, b.zzz);    // S1
    fq_mul(r, b.y, acc.zzz);     // S2
    fq_sub(pp, pp, u1);          // P
    fq_sub(r, r, s1);            // R
    if (fq_is_zero(pp)) {
        if (fq_is_zero(r))
            xyzz_dbl(acc);
        else
            xyzz_set_inf(acc);
        return;
    }
    fq_sqr(t, pp);               // PP
    fq_mul(pp, pp, t);           // PPP
    fq_mul(u1, u1, t);           // Q = U1*PP
    fq_mul(acc.zz, acc.zz, b.zz);
    fq_mul(acc.zz, acc.zz, t);   // ZZ3
    fq_mul(acc.zzz, acc.zzz, b.zzz);
    fq_mul(acc.zzz, acc.zzz, pp);// ZZZ3
    fq_sqr(t, r);
    fq_sub(t, t, pp);
    fq_sub(t, t, u1);
    fq_sub(acc.x, t, u1);        // X3
    fq_sub(u1, u1, acc.x);
    fq_mul(u1, r, u1);
    fq_mul(s1, s1, pp);
    fq_sub(acc.y, u1, s1);       // Y3
}

// XYZZ -> libff "special" projective {x, y, one} / {0, one, 0}
// (to_special / to_affine_coordinates, bw6_761_g1.cpp:58-77). One inversion.
__device__ __noinline__ void xyzz_to_special(Proj &out, const XYZZ &p)
{
    if (xyzz_is_inf(p)) {
        fq_set_zero(out.X);
        fq_set_one(out.Y);
        fq_set_zero(out.Z);
        return;
    }
    Fq zi, t;
    fq_inv(zi, p.zzz);           // 1/ZZZ
    fq_mul(out.Y, p.y, zi);      // y = Y/ZZZ
    fq_mul(t, zi, p.zz);         // ZZ/ZZZ = 1/Z
    fq_sqr(t, t);                // 1/ZZ
    fq_mul(out.X, p.x, t);       // x = X/ZZ
    fq_set_one(out.Z);
}

// XYZZ -> homogeneous projective without inversion: (X*ZZZ : Y*ZZ : ZZ*ZZZ).
__device__ __noinline__ void xyzz_to_proj(Proj &out, const XYZZ &p)
{
    if (xyzz_is_inf(p)) {
        fq_set_zero(out.X);
        fq_set_one(out.Y);
        fq_set_zero(out.Z);
        return;
    }
    fq_mul(out.X, p.x, p.zzz);
    fq_mul(out.Y, p.y, p.zz);
    fq_mul(out.Z, p.zz, p.zzz);
}

// libff projective {X,Y,Z} (any Z) -> XYZZ: x = X/Z = X*Z/Z^2, y = Y/Z = Y*Z^2/Z^3.
__device__ __noinline__ void xyzz_from_proj(XYZZ &r, const Proj &p)
{
    if (fq_is_zero(p.Z)) {
        xyzz_set_inf(r);
        return;
    }
    fq_sqr(r.zz, p.Z);
    fq_mul(r.zzz, r.zz, p.Z);
    fq_mul(r.x, p.X, p.Z);
    fq_mul(r.y, p.Y, r.zz);
}

} // namespace zecale
