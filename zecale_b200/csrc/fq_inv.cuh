// Plain modular inverse in Fq for sm_100a: batched-step binary GCD.
//
// Replaces the shift/subtract extended Euclid that served libff's Fp_model::invert
// (mpn_gcdext, fp.tcc:680-727) on the device. The 2*761-1 binary-GCD steps are run 30 at a
// time on 64-bit approximations of (a, b) (their low 30 bits and the 34 bits below the common
// top bit — the low bits decide parity, the top bits decide the comparison), branch-free, while
// recording the 2x2 integer transition matrix (f0 g0; f1 g1), |f|+|g| <= 2^30. The matrix is then
// applied to the full-width values with 32x32->64 multiply-adds:
//     a' = |a f0 + b g0| / 2^30        b' = |a f1 + b g1| / 2^30          (exact divisions)
//     u' = (u f0 + v g0) / 2^30 mod q  v' = (u f1 + v g1) / 2^30 mod q    (Montgomery-style)
// with invariants a = y u, b = y v (mod q). 52 rounds x 30 steps >= 1521 is the worst case; the loop
// ends as soon as a reaches 0 (b = gcd = 1, v = y^-1), after 34-38 rounds for random inputs. (T. Pornin, "Optimized binary GCD for modular
// inversion", 2020: the approximation never costs an extra step.)
// Cost: 52 x 240 multiply-adds of a 32-bit limb by a signed 31-bit factor plus the branch-free inner
// loops: ~140 k SASS instructions per warp as compiled (ncu), i.e. ~100 Montgomery products' worth
// of issue slots but mostly on the ALU pipe, and no data-dependent branch: a warp's 32 lanes invert 32
// different values in lock-step (2^19 inversions: 3.4 ms against 26 ms for the shift/subtract
// version, whose ~1500 iterations diverge four ways).
//
// The code is plain integer C++ so the same text compiles for the host (tests/test_inv_cpu.py
// checks it against Python's pow(y, -1, q)).
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define ZECALE_HD __host__ __device__ __forceinline__
#else
#define ZECALE_HD inline
#endif

namespace zecale {

static constexpr int INV_K = 30;       // steps per round
static constexpr int INV_ROUNDS = 52;  // 52*30 = 1560 >= 2*761-1

// (lo 30 bits) | (34 bits below the common top bit) << 30, for both a and b.
ZECALE_HD void inv_approx(const uint32_t (&a)[24], const uint32_t (&b)[24], uint64_t &ah, uint64_t &bh)
{
    // window = the three limbs ending at the highest limb where a|b is non-zero (index >= 2)
    uint32_t a2 = a[2], a1 = a[1], a0 = a[0], b2 = b[2], b1 = b[1], b0 = b[0];
    uint32_t big = a[2] | b[2];
#pragma unroll
    for (int i = 3; i < 24; i++) {
        const bool nz = (a[i] | b[i]) != 0;
        a0 = nz ? a[i - 2] : a0;
        a1 = nz ? a[i - 1] : a1;
        a2 = nz ? a[i] : a2;
        b0 = nz ? b[i - 2] : b0;
        b1 = nz ? b[i - 1] : b1;
        b2 = nz ? b[i] : b2;
        big |= a[i] | b[i];
    }
    if (big == 0) { // both fit in 64 bits: exact
        ah = (uint64_t)a[0] | ((uint64_t)a[1] << 32);
        bh = (uint64_t)b[0] | ((uint64_t)b[1] << 32);
        return;
    }
    const uint32_t top = a2 | b2;
    uint32_t s = 0; // leading zeros of top (top != 0)
    {
        uint32_t t = top;
        if (!(t & 0xffff0000u)) { s += 16; t <<= 16; }
        if (!(t & 0xff000000u)) { s += 8; t <<= 8; }
        if (!(t & 0xf0000000u)) { s += 4; t <<= 4; }
        if (!(t & 0xc0000000u)) { s += 2; t <<= 2; }
        if (!(t & 0x80000000u)) { s += 1; }
    }
    // top 64 bits of the 96-bit window, aligned to the common top bit
    const uint64_t ahi = ((uint64_t)a2 << 32) | a1, bhi = ((uint64_t)b2 << 32) | b1;
    const uint64_t at = s ? ((ahi << s) | ((uint64_t)a0 >> (32 - s))) : ahi;
    const uint64_t bt = s ? ((bhi << s) | ((uint64_t)b0 >> (32 - s))) : bhi;
    const uint64_t lo = (1ull << INV_K) - 1;
    ah = ((uint64_t)a[0] & lo) | ((at >> INV_K) << INV_K);
    bh = ((uint64_t)b[0] & lo) | ((bt >> INV_K) << INV_K);
}

// r = |x f + y g| >> 30 ; returns true if x f + y g was negative (r is the magnitude).
ZECALE_HD bool inv_lin_shift(uint32_t (&r)[24], const uint32_t (&x)[24], const uint32_t (&y)[24], int32_t f, int32_t g)
{
    uint32_t t[25];
    int64_t acc = 0;
#pragma unroll
    for (int i = 0; i < 24; i++) {
        acc += (int64_t)(uint64_t)x[i] * (int64_t)f + (int64_t)(uint64_t)y[i] * (int64_t)g;
        t[i] = (uint32_t)acc;
        acc >>= 32;
    }
    t[24] = (uint32_t)acc;
    const bool neg = acc < 0;
    // two's-complement negate of the 25-limb value when negative
    const uint32_t m = neg ? 0xffffffffu : 0u;
    uint64_t c = neg ? 1u : 0u;
#pragma unroll
    for (int i = 0; i < 25; i++) {
        c += (uint64_t)(t[i] ^ m);
        t[i] = (uint32_t)c;
        c >>= 32;
    }
#pragma unroll
    for (int i = 0; i < 24; i++) r[i] = (t[i] >> INV_K) | (t[i + 1] << (32 - INV_K));
    return neg;
}

// r = (x f + y g) / 2^30 mod q for 0 <= x, y <= q and |f| + |g| <= 2^30; r in [0, q).
ZECALE_HD void inv_lin_mod(uint32_t (&r)[24], const uint32_t (&x)[24], const uint32_t (&y)[24], int32_t f, int32_t g,
                           const uint32_t *P, uint32_t pinv30)
{
    // a negative factor is moved onto the operand: x f = (q - x) |f| (mod q)
    uint32_t xs[24], ys[24];
    {
        const bool nf = f < 0, ng = g < 0;
        int64_t cx = 0, cy = 0;
#pragma unroll
        for (int i = 0; i < 24; i++) {
            cx += (int64_t)P[i] - (int64_t)x[i];
            cy += (int64_t)P[i] - (int64_t)y[i];
            xs[i] = nf ? (uint32_t)cx : x[i];
            ys[i] = ng ? (uint32_t)cy : y[i];
            cx >>= 32;
            cy >>= 32;
        }
    }
    const uint32_t F = (uint32_t)(f < 0 ? -f : f), G = (uint32_t)(g < 0 ? -g : g);
    const uint32_t t0 = xs[0] * F + ys[0] * G;
    const uint32_t k = (t0 * pinv30) & ((1u << INV_K) - 1); // t + k q == 0 mod 2^30
    uint32_t t[25];
    uint64_t acc = 0;
#pragma unroll
    for (int i = 0; i < 24; i++) {
        // xs F + ys G <= 2^32 (F + G) <= 2^62, k q_i < 2^62, carry < 2^32: no overflow
        acc += (uint64_t)xs[i] * F + (uint64_t)ys[i] * G + (uint64_t)P[i] * k;
        t[i] = (uint32_t)acc;
        acc >>= 32;
    }
    t[24] = (uint32_t)acc;
    uint32_t v[24];
#pragma unroll
    for (int i = 0; i < 24; i++) v[i] = (t[i] >> INV_K) | (t[i + 1] << (32 - INV_K));
    // v < 3 q: subtract q up to twice
#pragma unroll
    for (int pass = 0; pass < 2; pass++) {
        uint32_t d[24];
        int64_t c = 0;
#pragma unroll
        for (int i = 0; i < 24; i++) {
            c += (int64_t)v[i] - (int64_t)P[i];
            d[i] = (uint32_t)c;
            c >>= 32;
        }
        const bool ge = c == 0; // no borrow
#pragma unroll
        for (int i = 0; i < 24; i++) v[i] = ge ? d[i] : v[i];
    }
#pragma unroll
    for (int i = 0; i < 24; i++) r[i] = v[i];
}

// r = y^-1 mod q for a plain integer 0 < y < q (q odd, 761 bits, limbs in P; pinv32 = -q^-1 mod 2^32).
// y == 0 returns 0.
ZECALE_HD void fq_inv_bingcd(uint32_t (&r)[24], const uint32_t (&y)[24], const uint32_t *P, uint32_t pinv32)
{
    uint32_t a[24], b[24], u[24], v[24];
#pragma unroll
    for (int i = 0; i < 24; i++) {
        a[i] = y[i];
        b[i] = P[i];
        u[i] = i == 0 ? 1u : 0u;
        v[i] = 0;
    }
    const uint32_t pinv30 = pinv32 & ((1u << INV_K) - 1);
#pragma unroll 1
    for (int round = 0; round < INV_ROUNDS; round++) {
        // a == 0: b = gcd = 1 and v = y^-1 are final. Random inputs get here after 34-38 rounds; the
        // lanes of a warp leave one by one and the warp runs as long as its slowest lane.
        uint32_t a_or = 0;
#pragma unroll
        for (int i = 0; i < 24; i++) a_or |= a[i];
        if (a_or == 0) break;
        uint64_t ah, bh;
        inv_approx(a, b, ah, bh);
        int32_t f0 = 1, g0 = 0, f1 = 0, g1 = 1;
#pragma unroll 2
        for (int i = 0; i < INV_K; i++) {
            const bool odd = (ah & 1u) != 0;
            const bool sw = odd && ah < bh;
            const uint64_t ta = sw ? bh : ah, tb = sw ? ah : bh;
            const int32_t tf0 = sw ? f1 : f0, tf1 = sw ? f0 : f1, tg0 = sw ? g1 : g0, tg1 = sw ? g0 : g1;
            ah = (ta - (odd ? tb : 0)) >> 1;
            bh = tb;
            f0 = tf0 - (odd ? tf1 : 0);
            g0 = tg0 - (odd ? tg1 : 0);
            f1 = (int32_t)((uint32_t)tf1 << 1);
            g1 = (int32_t)((uint32_t)tg1 << 1);
        }
        uint32_t na[24], nb[24];
        const bool nega = inv_lin_shift(na, a, b, f0, g0);
        const bool negb = inv_lin_shift(nb, a, b, f1, g1);
        if (nega) {
            f0 = -f0;
            g0 = -g0;
        }
        if (negb) {
            f1 = -f1;
            g1 = -g1;
        }
        uint32_t nu[24], nv[24];
        inv_lin_mod(nu, u, v, f0, g0, P, pinv30);
        inv_lin_mod(nv, u, v, f1, g1, P, pinv30);
#pragma unroll
        for (int i = 0; i < 24; i++) {
            a[i] = na[i];
            b[i] = nb[i];
            u[i] = nu[i];
            v[i] = nv[i];
        }
    }
    // y == 0: a stays 0, b stays q, v stays 0
#pragma unroll
    for (int i = 0; i < 24; i++) r[i] = v[i];
}

} // namespace zecale
