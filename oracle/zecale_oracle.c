/* zecale_oracle.c — CPU restatement ("port") of the reference algorithms on the BW6-761
 * Groth16 prove hot path. TEST INFRASTRUCTURE ONLY: may be used by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline leg as the CHECKER — never by
 * zecale_b200/ (the product path has no CPU fallback).
 *
 * Parity status: PINNED. Every entry point below is checked bit-for-bit against
 *   (a) oracle/_ref (the unmodified libff/libfqfft/libsnark compiled from /root/reference by
 *       oracle/Makefile) wherever that library exists (tests/test_oracle_cpu.py), and
 *   (b) the golden fixtures under tests/golden/ that tests/golden/make_golden.py produced
 *       from oracle/_ref in the build container (these travel to the GPU box).
 * The reference repository ships no value-level golden vectors for this path.
 *
 * Reference abbreviations: LF = depends/zeth/depends/libsnark/depends/libff/libff,
 * LQ = .../depends/libfqfft/libfqfft, LS = depends/zeth/depends/libsnark/libsnark.
 *
 * Plain C, no GMP: limbs are uint64_t little-endian exactly like libff's bigint<n>
 * (LF/algebra/fields/bigint.hpp:28-65); products use unsigned __int128.
 */
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "zecale_oracle_consts.h"

typedef unsigned __int128 u128;
typedef uint64_t u64;

/* A prime field in Montgomery form, R = 2^(64 n) (LF/algebra/fields/fp.hpp:38-160). */
typedef struct {
    int n;
    const u64 *mod, *one, *r2, *r3, *pm2;
    u64 inv; /* -mod^-1 mod 2^64 */
    int bits;
} field_t;

static const field_t FQ = {12, ZQ_MOD, ZQ_ONE, ZQ_R2, ZQ_R3, ZQ_PM2, ZQ_INV, ZQ_BITS};
static const field_t FR = {6, ZR_MOD, ZR_ONE, ZR_R2, ZR_R3, ZR_PM2, ZR_INV, ZR_BITS};

#define MAXN 12

static int big_is_zero(const u64 *a, int n)
{
    u64 acc = 0;
    for (int i = 0; i < n; i++) acc |= a[i];
    return acc == 0;
}

static int big_cmp(const u64 *a, const u64 *b, int n)
{
    for (int i = n - 1; i >= 0; i--) {
        if (a[i] != b[i]) return a[i] > b[i] ? 1 : -1;
    }
    return 0;
}

static u64 big_add(u64 *r, const u64 *a, const u64 *b, int n)
{
    u128 c = 0;
    for (int i = 0; i < n; i++) {
        c += (u128)a[i] + b[i];
        r[i] = (u64)c;
        c >>= 64;
    }
    return (u64)c;
}

static u64 big_sub(u64 *r, const u64 *a, const u64 *b, int n)
{
    u64 borrow = 0;
    for (int i = 0; i < n; i++) {
        u128 d = (u128)a[i] - b[i] - borrow;
        r[i] = (u64)d;
        borrow = (u64)(d >> 64) & 1;
    }
    return borrow;
}

/* r = a + b mod p   (fp.tcc:351-465, generic branch :448-462: add, then subtract the modulus
 * if there was a carry or the sum is >= modulus). */
static void fp_add(const field_t *F, u64 *r, const u64 *a, const u64 *b)
{
    u64 t[MAXN];
    u64 carry = big_add(t, a, b, F->n);
    if (carry || big_cmp(t, F->mod, F->n) >= 0) big_sub(t, t, F->mod, F->n);
    memcpy(r, t, 8 * F->n);
}

/* r = a - b mod p   (fp.tcc:468-547, generic :528-545: if a < b add the modulus first). */
static void fp_sub(const field_t *F, u64 *r, const u64 *a, const u64 *b)
{
    u64 t[MAXN];
    if (big_sub(t, a, b, F->n)) big_add(t, t, F->mod, F->n);
    memcpy(r, t, 8 * F->n);
}

/* r = -a mod p, 0 stays 0   (fp.tcc:617-630). */
static void fp_neg(const field_t *F, u64 *r, const u64 *a)
{
    if (big_is_zero(a, F->n)) {
        memset(r, 0, 8 * F->n);
        return;
    }
    big_sub(r, F->mod, a, F->n);
}

/* Montgomery product r = a b R^-1 mod p, fully reduced to [0, p)   (mul_reduce, fp.tcc:51-228;
 * the generic branch :204-227 forms the 2n-limb product and then runs n rounds of
 * "k = t[i] * inv; t += k * mod << 64 i", final conditional subtract :219-224). */
static void fp_mul(const field_t *F, u64 *r, const u64 *a, const u64 *b)
{
    const int n = F->n;
    u64 t[2 * MAXN + 1];
    memset(t, 0, sizeof t);
    for (int i = 0; i < n; i++) {
        u128 c = 0;
        for (int j = 0; j < n; j++) {
            c += (u128)a[i] * b[j] + t[i + j];
            t[i + j] = (u64)c;
            c >>= 64;
        }
        t[i + n] = (u64)c;
    }
    for (int i = 0; i < n; i++) {
        const u64 k = t[i] * F->inv;
        u128 c = 0;
        for (int j = 0; j < n; j++) {
            c += (u128)k * F->mod[j] + t[i + j];
            t[i + j] = (u64)c;
            c >>= 64;
        }
        for (int j = i + n; c && j <= 2 * n; j++) {
            c += t[j];
            t[j] = (u64)c;
            c >>= 64;
        }
    }
    if (t[2 * n] || big_cmp(t + n, F->mod, n) >= 0) big_sub(t + n, t + n, F->mod, n);
    memcpy(r, t + n, 8 * n);
}

/* squared() is a plain product for these limb counts (fp.tcc:633-677, :673-676). */
static void fp_sqr(const field_t *F, u64 *r, const u64 *a) { fp_mul(F, r, a, a); }

/* as_bigint(): Montgomery -> plain integer, i.e. a product with 1 (fp.tcc:270-281). */
static void fp_from_mont(const field_t *F, u64 *r, const u64 *a)
{
    u64 one[MAXN] = {1};
    fp_mul(F, r, a, one);
}

/* Constructor from a plain bigint: multiply by R^2 (fp.tcc:231-235). */
static void fp_to_mont(const field_t *F, u64 *r, const u64 *a) { fp_mul(F, r, a, F->r2); }

/* a^e for a plain multi-limb exponent, square-and-multiply from the top bit
 * (LF/algebra/exponentiation/exponentiation.tcc: power()). */
static void fp_pow(const field_t *F, u64 *r, const u64 *a, const u64 *e, int elimbs)
{
    u64 acc[MAXN], base[MAXN];
    memcpy(acc, F->one, 8 * F->n);
    memcpy(base, a, 8 * F->n);
    int started = 0;
    for (int bit = 64 * elimbs - 1; bit >= 0; bit--) {
        if (started) fp_sqr(F, acc, acc);
        if ((e[bit >> 6] >> (bit & 63)) & 1) {
            started = 1;
            fp_mul(F, acc, acc, base);
        }
    }
    memcpy(r, acc, 8 * F->n);
}

/* invert(): the reference runs mpn_gcdext on the Montgomery representation and multiplies by
 * R^3 (fp.tcc:680-727). The inverse in a prime field is unique, so a^(p-2) gives the same
 * canonical limbs; 0 is mapped to 0 here (the reference asserts non-zero). */
static void fp_inv(const field_t *F, u64 *r, const u64 *a) { fp_pow(F, r, a, F->pm2, F->n); }

static int fp_eq(const field_t *F, const u64 *a, const u64 *b) { return memcmp(a, b, 8 * F->n) == 0; }

/* op codes shared with the GPU test hooks (include/zecale_gpu.h):
 * 0 add, 1 sub, 2 mul, 3 sqr, 4 inverse, 5 neg, 6 from-Montgomery, 7 to-Montgomery */
static void field_op(const field_t *F, int op, const u64 *a, const u64 *b, u64 *o, size_t cnt)
{
    const int n = F->n;
    for (size_t i = 0; i < cnt; i++) {
        const u64 *x = a + i * n, *y = b ? b + i * n : x;
        u64 *z = o + i * n;
        switch (op) {
        case 0: fp_add(F, z, x, y); break;
        case 1: fp_sub(F, z, x, y); break;
        case 2: fp_mul(F, z, x, y); break;
        case 3: fp_sqr(F, z, x); break;
        case 4: fp_inv(F, z, x); break;
        case 5: fp_neg(F, z, x); break;
        case 6: fp_from_mont(F, z, x); break;
        case 7: fp_to_mont(F, z, x); break;
        default: memset(z, 0, 8 * n);
        }
    }
}

void zor_fq_op(int op, const void *a, const void *b, void *o, size_t n) { field_op(&FQ, op, a, b, o, n); }
void zor_fr_op(int op, const void *a, const void *b, void *o, size_t n) { field_op(&FR, op, a, b, o, n); }

/* ------------------------------------------------------------------ groups
 * bw6_761_G1 / bw6_761_G2: short Weierstrass curves with a = 0 over the SAME Fq, homogeneous
 * projective coordinates {X, Y, Z}, 288 bytes (LF/algebra/curves/bw6_761/bw6_761_g1.hpp:14-91,
 * bw6_761_g2.hpp). The curve constant b never enters add/dbl, so one set of routines serves both
 * groups (bw6_761_g2.cpp:116-364 repeats bw6_761_g1.cpp:111-358 verbatim over bw6_761_Fq). */
typedef struct {
    u64 X[12], Y[12], Z[12];
} point_t;

#define QM(r, a, b) fp_mul(&FQ, r, a, b)
#define QS(r, a) fp_sqr(&FQ, r, a)
#define QA(r, a, b) fp_add(&FQ, r, a, b)
#define QD(r, a, b) fp_sub(&FQ, r, a, b)

/* is_zero: X == 0 && Z == 0 (bw6_761_g1.cpp:79-82); zero() = (0, 1, 0) (bw6_761_init.cpp:285-286). */
static int pt_is_zero(const point_t *p) { return big_is_zero(p->X, 12) && big_is_zero(p->Z, 12); }

static void pt_set_zero(point_t *p)
{
    memset(p, 0, sizeof *p);
    memcpy(p->Y, ZQ_ONE, 96);
}

/* dbl(): dbl-2007-bl with a = 0 (bw6_761_g1.cpp:318-358). */
static void pt_dbl(point_t *r, const point_t *p)
{
    if (pt_is_zero(p)) {
        *r = *p;
        return;
    }
    u64 XX[12], w[12], s[12], ss[12], sss[12], R[12], RR[12], B[12], h[12], t[12];
    QS(XX, p->X);
    QA(w, XX, XX);
    QA(w, w, XX);
    QM(s, p->Y, p->Z);
    QA(s, s, s);
    QS(ss, s);
    QM(sss, s, ss);
    QM(R, p->Y, s);
    QS(RR, R);
    QA(t, p->X, R);
    QS(B, t);
    QD(B, B, XX);
    QD(B, B, RR);
    QS(h, w);
    QA(t, B, B);
    QD(h, h, t);
    point_t o;
    QM(o.X, h, s);
    QD(t, B, h);
    QM(o.Y, w, t);
    QA(t, RR, RR);
    QD(o.Y, o.Y, t);
    memcpy(o.Z, sss, 96);
    *r = o;
}

/* The shared tail of add-1998-cmo-2 / madd-1998-cmo once u, v, X1Z2, Y1Z2, Z1Z2 are known
 * (bw6_761_g1.cpp:168-193, 241-257, 303-315). */
static void pt_add_tail(point_t *r, const u64 *u, const u64 *v, const u64 *X1Z2, const u64 *Y1Z2, const u64 *Z1Z2)
{
    u64 uu[12], vv[12], vvv[12], R[12], A[12], t[12];
    QS(uu, u);
    QS(vv, v);
    QM(vvv, v, vv);
    QM(R, vv, X1Z2);
    QM(A, uu, Z1Z2);
    QA(t, R, R);
    QA(t, t, vvv);
    QD(A, A, t);
    point_t o;
    QM(o.X, v, A);
    QD(t, R, A);
    QM(o.Y, u, t);
    QM(t, vvv, Y1Z2);
    QD(o.Y, o.Y, t);
    QM(o.Z, vvv, Z1Z2);
    *r = o;
}

/* operator+ : general addition with the doubling fallback (bw6_761_g1.cpp:111-194). */
static void pt_add(point_t *r, const point_t *p, const point_t *q)
{
    if (pt_is_zero(p)) {
        *r = *q;
        return;
    }
    if (pt_is_zero(q)) {
        *r = *p;
        return;
    }
    u64 X1Z2[12], X2Z1[12], Y1Z2[12], Y2Z1[12], Z1Z2[12], u[12], v[12];
    QM(X1Z2, p->X, q->Z);
    QM(X2Z1, p->Z, q->X);
    QM(Y1Z2, p->Y, q->Z);
    QM(Y2Z1, p->Z, q->Y);
    if (fp_eq(&FQ, X1Z2, X2Z1) && fp_eq(&FQ, Y1Z2, Y2Z1)) {
        pt_dbl(r, p);
        return;
    }
    QM(Z1Z2, p->Z, q->Z);
    QD(u, Y2Z1, Y1Z2);
    QD(v, X2Z1, X1Z2);
    pt_add_tail(r, u, v, X1Z2, Y1Z2, Z1Z2);
}

/* mixed_add: q is in special form, Z2 = 1 or q = zero (bw6_761_g1.cpp:263-316). */
static void pt_mixed_add(point_t *r, const point_t *p, const point_t *q)
{
    if (pt_is_zero(p)) {
        *r = *q;
        return;
    }
    if (pt_is_zero(q)) {
        *r = *p;
        return;
    }
    u64 X2Z1[12], Y2Z1[12], u[12], v[12];
    QM(X2Z1, p->Z, q->X);
    QM(Y2Z1, p->Z, q->Y);
    if (fp_eq(&FQ, p->X, X2Z1) && fp_eq(&FQ, p->Y, Y2Z1)) {
        pt_dbl(r, p);
        return;
    }
    QD(u, Y2Z1, p->Y);
    QD(v, X2Z1, p->X);
    pt_add_tail(r, u, v, p->X, p->Y, p->Z);
}

/* operator- (unary): (X, -Y, Z) (bw6_761_g1.cpp:196-199). */
static void pt_neg(point_t *r, const point_t *p)
{
    point_t o = *p;
    fp_neg(&FQ, o.Y, p->Y);
    *r = o;
}

/* to_special() = to_affine_coordinates(): zero -> (0, 1, 0), else (X/Z, Y/Z, 1)
 * (bw6_761_g1.cpp:58-77). */
static void pt_to_special(point_t *r, const point_t *p)
{
    if (pt_is_zero(p)) {
        pt_set_zero(r);
        return;
    }
    u64 zi[12];
    point_t o;
    fp_inv(&FQ, zi, p->Z);
    QM(o.X, p->X, zi);
    QM(o.Y, p->Y, zi);
    memcpy(o.Z, ZQ_ONE, 96);
    *r = o;
}

/* op: 0 operator+, 1 mixed_add, 2 dbl, 3 to_special, 4 negate. Results of 0-2 are returned in
 * special form too so that they can be compared byte-for-byte (projective triples are not unique). */
void zor_group_op(int op, const void *a_, const void *b_, void *o_, size_t n)
{
    const point_t *a = a_, *b = b_ ? b_ : a_;
    point_t *o = o_;
    for (size_t i = 0; i < n; i++) {
        point_t t;
        switch (op) {
        case 0: pt_add(&t, a + i, b + i); break;
        case 1: pt_mixed_add(&t, a + i, b + i); break;
        case 2: pt_dbl(&t, a + i); break;
        case 4: pt_neg(&t, a + i); break;
        default: t = a[i];
        }
        pt_to_special(o + i, &t);
    }
}

/* ------------------------------------------------------------------ signed windows
 * field_get_digit / field_get_signed_digit (LF/algebra/fields/field_utils.tcc:50-115): the c-bit
 * window `idx` of a plain 6-limb scalar, recoded to [-2^(c-1), 2^(c-1)) with the carry rippling
 * up from window 0. */
static u64 scalar_window(const u64 *v, unsigned c, unsigned idx)
{
    const unsigned start = c * idx;
    if (start >= 384) return 0;
    const unsigned lo = start >> 6, sh = start & 63;
    u64 w = v[lo] >> sh;
    if (sh + c > 64 && lo + 1 < 6) w |= v[lo + 1] << (64 - sh);
    return w & ((1ull << c) - 1);
}

long zor_fr_signed_digit(const void *plain, size_t c, size_t idx)
{
    const u64 *v = plain;
    const u64 half = 1ull << (c - 1), full = 1ull << c;
    u64 carry = 0, digit = 0;
    int overflow = 0;
    for (size_t i = 0; i <= idx; i++) {
        digit = scalar_window(v, (unsigned)c, (unsigned)i) + carry;
        overflow = (digit & full) != 0;      /* window was all ones and a carry came in */
        carry = overflow | ((digit & half) != 0);
    }
    if (overflow) return 0;
    return (digit & half) ? (long)digit - (long)full : (long)digit;
}

/* pippenger_optimal_c / bdlo12_signed_optimal_c (multiexp.tcc:35-40, 637-641). */
static unsigned log2_ceil(size_t n)
{ /* libff::log2 = ceil(log2 n) (LF/common/utils.cpp) */
    unsigned r = (n & (n - 1)) == 0 ? 0 : 1;
    while (n > 1) {
        n >>= 1;
        r++;
    }
    return r;
}

size_t zor_optimal_c(size_t n)
{
    const size_t l = log2_ceil(n);
    return l - (l / 3 - 2) + 1;
}

static unsigned scalar_bits(const u64 *v)
{
    for (int i = 5; i >= 0; i--)
        if (v[i]) return 64 * i + 64 - (unsigned)__builtin_clzll(v[i]);
    return 0;
}

/* ------------------------------------------------------------------ MSM
 * multi_exp_implementation<G, Fr, BDLO12_signed, special>::multi_exp_inner (multiexp.tcc:563-632):
 * scalars de-Montgomeryised once; rounds = ceil((max_bits + 2) / c); one window at a time from the
 * top, c doublings between windows, 2^(c-1) buckets reused; bucket fill by mixed addition of
 * +/- base (:45-81), bucket fold by the running-sum trick (:90-125). */
static void msm_inner(point_t *out, const uint8_t *bases, size_t bstride, const uint8_t *scalars, size_t sstride,
                      const uint64_t *index, size_t n)
{
    if (n == 0) {
        pt_set_zero(out);
        return;
    }
    const unsigned c = (unsigned)zor_optimal_c(n);
    u64 *plain = malloc(n * 48);
    unsigned bits = 0;
    for (size_t i = 0; i < n; i++) {
        const size_t si = index ? (size_t)index[i] : i;
        fp_from_mont(&FR, plain + 6 * i, (const u64 *)(scalars + si * sstride));
        const unsigned b = scalar_bits(plain + 6 * i);
        if (b > bits) bits = b;
    }
    const unsigned rounds = (bits + 2 + c - 1) / c;
    const size_t nb = (size_t)1 << (c - 1);
    point_t *bucket = malloc(nb * sizeof(point_t));
    uint8_t *hit = malloc(nb);
    point_t result;
    pt_set_zero(&result);
    for (unsigned round = 0; round < rounds; round++) {
        const unsigned w = rounds - 1 - round;
        if (round)
            for (unsigned k = 0; k < c; k++) pt_dbl(&result, &result);
        memset(hit, 0, nb);
        size_t nonzero = 0;
        for (size_t i = 0; i < n; i++) {
            const long d = zor_fr_signed_digit(plain + 6 * i, c, w);
            if (d == 0) continue;
            point_t b = *(const point_t *)(bases + i * bstride);
            size_t k;
            if (d < 0) {
                pt_neg(&b, &b);
                k = (size_t)(-d) - 1;
            } else {
                k = (size_t)d - 1;
            }
            if (hit[k]) {
                pt_mixed_add(&bucket[k], &bucket[k], &b);
            } else {
                bucket[k] = b;
                hit[k] = 1;
            }
            nonzero++;
        }
        if (!nonzero) continue; /* the round contributes zero() */
        size_t i = nb - 1;
        while (!hit[i]) i--;
        point_t acc = bucket[i], sum = bucket[i];
        while (i > 0) {
            i--;
            if (hit[i]) pt_add(&acc, &acc, &bucket[i]);
            pt_add(&sum, &sum, &acc);
        }
        pt_add(&result, &result, &sum);
    }
    free(hit);
    free(bucket);
    free(plain);
    *out = result;
}

/* multi_exp (multiexp.tcc:648-688): `chunks` equal ranges (the last takes the remainder), one
 * multi_exp_inner each (OpenMP), partial results summed in order. chunks <= 1 or n < chunks:
 * a single inner call. Result in special form. index (nullable) is the sparse walk of
 * kc_multi_exp_with_mixed_addition (LS/knowledge_commitment/kc_multiexp.tcc:63-89): the scalar of
 * base i is scalars[index[i]]. */
void zor_msm(const void *bases, size_t bstride, const void *scalars, size_t sstride, const uint64_t *index, size_t n,
             size_t chunks, void *out)
{
    point_t total;
    if (n < chunks || chunks <= 1) {
        msm_inner(&total, bases, bstride, scalars, sstride, index, n);
    } else {
        const size_t one = n / chunks;
        point_t *part = malloc(chunks * sizeof(point_t));
#pragma omp parallel for schedule(dynamic, 1)
        for (size_t k = 0; k < chunks; k++) {
            const size_t lo = k * one, hi = k == chunks - 1 ? n : (k + 1) * one;
            /* dense: scalar i belongs to base i; sparse: the index list carries the offset */
            const uint8_t *sc = index ? (const uint8_t *)scalars : (const uint8_t *)scalars + lo * sstride;
            msm_inner(&part[k], (const uint8_t *)bases + lo * bstride, bstride, sc, sstride, index ? index + lo : NULL,
                      hi - lo);
        }
        pt_set_zero(&total);
        for (size_t k = 0; k < chunks; k++) pt_add(&total, &total, &part[k]);
        free(part);
    }
    pt_to_special(out, &total);
}

/* ------------------------------------------------------------------ NTT
 * get_root_of_unity (LF/algebra/fields/field_utils.tcc:190-206): omega_m =
 * root_of_unity^(2^(s - log m)) by repeated squaring. */
static void fr_omega(u64 *w, unsigned log_m)
{
    memcpy(w, ZR_ROOT_2S, 48);
    for (unsigned i = log_m; i < ZR_TWO_ADICITY; i++) fp_sqr(&FR, w, w);
}

static size_t bitrev(size_t x, unsigned bits)
{
    size_t r = 0;
    for (unsigned i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}

/* _basic_serial_radix2_FFT (LQ/evaluation_domain/domains/basic_radix2_domain_aux.tcc:44-79):
 * bit-reversal swap, then log m DIT stages with a running twiddle; no scaling. */
static void fr_fft(u64 *a, unsigned log_m, const u64 *omega)
{
    const size_t n = (size_t)1 << log_m;
    for (size_t k = 0; k < n; k++) {
        const size_t rk = bitrev(k, log_m);
        if (k < rk) {
            u64 t[6];
            memcpy(t, a + 6 * k, 48);
            memcpy(a + 6 * k, a + 6 * rk, 48);
            memcpy(a + 6 * rk, t, 48);
        }
    }
    size_t m = 1;
    for (unsigned s = 1; s <= log_m; s++) {
        u64 w_m[6], e[1] = {n / (2 * m)};
        fp_pow(&FR, w_m, omega, e, 1);
        for (size_t k = 0; k < n; k += 2 * m) {
            u64 w[6];
            memcpy(w, ZR_ONE, 48);
            for (size_t j = 0; j < m; j++) {
                u64 t[6], *lo = a + 6 * (k + j), *hi = a + 6 * (k + j + m);
                fp_mul(&FR, t, w, hi);
                fp_sub(&FR, hi, lo, t);
                fp_add(&FR, lo, lo, t);
                fp_mul(&FR, w, w, w_m);
            }
        }
        m *= 2;
    }
}

/* _multiply_by_coset (basic_radix2_domain_aux.tcc:174-183): a[i] *= g^i for i >= 1. */
static void fr_mul_coset(u64 *a, size_t n, const u64 *g)
{
    u64 u[6];
    memcpy(u, g, 48);
    for (size_t i = 1; i < n; i++) {
        fp_mul(&FR, a + 6 * i, a + 6 * i, u);
        fp_mul(&FR, u, u, g);
    }
}

/* basic_radix2_domain<Fr>::FFT / iFFT / cosetFFT / icosetFFT (basic_radix2_domain.tcc:55-88):
 * flags bit0 = inverse (omega^-1, then * m^-1), bit1 = coset (forward: * g^i first; inverse:
 * * g^-i last). In place, natural order, Montgomery form. */
int zor_ntt(void *a_, unsigned log_m, unsigned flags, const void *coset_g)
{
    if (log_m < 1 || log_m > ZR_TWO_ADICITY) return 1;
    u64 *a = a_;
    const size_t m = (size_t)1 << log_m;
    u64 omega[6];
    fr_omega(omega, log_m);
    if (!(flags & 1)) {
        if (flags & 2) fr_mul_coset(a, m, coset_g);
        fr_fft(a, log_m, omega);
        return 0;
    }
    u64 oinv[6], minv[6], mm[6] = {m};
    fp_inv(&FR, oinv, omega);
    fr_fft(a, log_m, oinv);
    fp_to_mont(&FR, mm, mm);
    fp_inv(&FR, minv, mm);
    for (size_t i = 0; i < m; i++) fp_mul(&FR, a + 6 * i, a + 6 * i, minv);
    if (flags & 2) {
        u64 ginv[6];
        fp_inv(&FR, ginv, coset_g);
        fr_mul_coset(a, m, ginv);
    }
    return 0;
}

/* The transform part of r1cs_to_qap_witness_map for d1 = d2 = d3 = 0
 * (LS/reductions/r1cs_to_qap/r1cs_to_qap.tcc:252-330): aA, aB, aC are the evaluations on the
 * domain; iFFT + cosetFFT each, H_tmp = aA∘aB - aC on the coset, divide_by_Z_on_coset (the
 * vanishing polynomial is the constant g^m - 1 there, basic_radix2_domain.tcc:103-126),
 * icosetFFT. out: m + 1 coefficients, out[m] = 0. Inputs are not modified. */
int zor_qap_h(const void *aA, const void *aB, const void *aC, unsigned log_m, const void *g, void *out)
{
    const size_t m = (size_t)1 << log_m;
    u64 *v[3];
    const void *src[3] = {aA, aB, aC};
    for (int k = 0; k < 3; k++) {
        v[k] = malloc(m * 48);
        memcpy(v[k], src[k], m * 48);
        if (zor_ntt(v[k], log_m, 1, NULL) || zor_ntt(v[k], log_m, 2, g)) return 1;
    }
    u64 z[6], zinv[6], e[1] = {m};
    fp_pow(&FR, z, g, e, 1);
    fp_sub(&FR, z, z, ZR_ONE);
    fp_inv(&FR, zinv, z);
    u64 *h = out;
    for (size_t i = 0; i < m; i++) {
        u64 t[6];
        fp_mul(&FR, t, v[0] + 6 * i, v[1] + 6 * i);
        fp_sub(&FR, t, t, v[2] + 6 * i);
        fp_mul(&FR, h + 6 * i, t, zinv);
    }
    memset(h + 6 * m, 0, 48);
    for (int k = 0; k < 3; k++) free(v[k]);
    return zor_ntt(h, log_m, 3, g);
}
