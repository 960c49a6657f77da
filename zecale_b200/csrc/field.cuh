// BW6-761 field layer for sm_100a: Fq (761 bit, 24 x u32) and Fr (377 bit, 12 x u32).
//
// Replaces libff's Fp_model<12, bw6_761_modulus_q> / Fp_model<6, bw6_761_modulus_r>
// (libff/algebra/fields/fp.hpp:38-160, fp.tcc:204-227 mul_reduce, :448-462 add,
// :528-545 sub, :270-281 as_bigint, :680-727 invert). Same Montgomery radix
// (R = 2^768 / 2^384) and fully reduced representatives, so a device element is
// byte-identical to the reference's mont_repr on a little-endian host.
//
// Layout decision: a 761-bit element does not fit the register file next to the
// temporaries of a point addition (10 live Fq = 240 registers), so elements live in
// memory (local / shared / global, reached through generic pointers) and only the
// arithmetic kernels — fq_mul & co., one out-of-line copy each per module — pull
// operands into registers. 72 loads/stores per 1176 IMAD.WIDE is 6 % overhead and it
// keeps the SASS of a whole MSM kernel under the 32 KB L1.5 instruction cache.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "field_asm.cuh"
#include "field_consts.cuh"

namespace zecale {

struct alignas(16) Fq {
    uint32_t l[24];
};
struct alignas(16) Fr {
    uint32_t l[12];
};
static_assert(sizeof(Fq) == 96 && sizeof(Fr) == 48, "field element layout");

// ---------------------------------------------------------------- Fq (memory form)
__device__ __noinline__ void fq_mul(Fq &r, const Fq &a, const Fq &b)
{
    uint32_t x[24], y[24], z[24];
#pragma unroll
    for (int i = 0; i < 24; i++) {
        x[i] = a.l[i];
        y[i] = b.l[i];
    }
    fq_mul_regs(z, x, y);
#pragma unroll
    for (int i = 0; i < 24; i++) r.l[i] = z[i];
}

__device__ __noinline__ void fq_sqr(Fq &r, const Fq &a)
{
    uint32_t x[24], z[24];
#pragma unroll
    for (int i = 0; i < 24; i++) x[i] = a.l[i];
    fq_sqr_regs(z, x);
#pragma unroll
    for (int i = 0; i < 24; i++) r.l[i] = z[i];
}

__device__ __noinline__ void fq_add(Fq &r, const Fq &a, const Fq &b)
{
    uint32_t x[24], y[24], z[24];
#pragma unroll
    for (int i = 0; i < 24; i++) {
        x[i] = a.l[i];
        y[i] = b.l[i];
    }
    fq_add_regs(z, x, y);
#pragma unroll
    for (int i = 0; i < 24; i++) r.l[i] = z[i];
}

__device__ __noinline__ void fq_sub(Fq &r, const Fq &a, const Fq &b)
{
    uint32_t x[24], y[24], z[24];
#pragma unroll
    for (int i = 0; i < 24; i++) {
        x[i] = a.l[i];
        y[i] = b.l[i];
    }
    fq_sub_regs(z, x, y);
#pragma unroll
    for (int i = 0; i < 24; i++) r.l[i] = z[i];
}

__device__ __forceinline__ void fq_copy(Fq &r, const Fq &a)
{
    const uint4 *s = reinterpret_cast<const uint4 *>(a.l);
    uint4 *d = reinterpret_cast<uint4 *>(r.l);
#pragma unroll
    for (int i = 0; i < 6; i++) d[i] = s[i];
}

__device__ __forceinline__ void fq_set_zero(Fq &r)
{
    uint4 *d = reinterpret_cast<uint4 *>(r.l);
#pragma unroll
    for (int i = 0; i < 6; i++) d[i] = make_uint4(0, 0, 0, 0);
}

__device__ __forceinline__ void fq_set_one(Fq &r)
{
#pragma unroll
    for (int i = 0; i < 24; i++) r.l[i] = FQ_ONE[i];
}

__device__ __forceinline__ bool fq_is_zero(const Fq &a)
{
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 24; i++) acc |= a.l[i];
    return acc == 0;
}

__device__ __forceinline__ bool fq_eq(const Fq &a, const Fq &b)
{
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 24; i++) acc |= a.l[i] ^ b.l[i];
    return acc == 0;
}

// r = -a mod q (0 stays 0), cf. fp.tcc:617-630
__device__ __noinline__ void fq_neg(Fq &r, const Fq &a)
{
    Fq z;
    fq_set_zero(z);
    fq_sub(r, z, a);
}

__device__ __forceinline__ void fq_dbl(Fq &r, const Fq &a) { fq_add(r, a, a); }

// r = a^(q-2) (Fermat). 0 -> 0. ~761 squarings + ~380 products; only used to
// normalise the single output point of an MSM (libff uses mpn_gcdext, fp.tcc:680-727;
// the inverse is unique so the result is identical).
__device__ __noinline__ void fq_inv(Fq &r, const Fq &a)
{
    Fq acc, base;
    fq_set_one(acc);
    fq_copy(base, a);
    for (int bit = FQ_BITS - 1; bit >= 0; bit--) {
        fq_sqr(acc, acc);
        if ((FQ_PM2[bit >> 5] >> (bit & 31)) & 1u) fq_mul(acc, acc, base);
    }
    fq_copy(r, acc);
}

// ---------------------------------------------------------------- Fr (register form)
// Fr elements are small enough (12 registers) to be handled by value.
__device__ __forceinline__ void fr_load(uint32_t (&r)[12], const Fr *p)
{
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
#pragma unroll
    for (int i = 0; i < 3; i++) {
        uint4 v = s[i];
        r[4 * i] = v.x;
        r[4 * i + 1] = v.y;
        r[4 * i + 2] = v.z;
        r[4 * i + 3] = v.w;
    }
}

__device__ __forceinline__ void fr_store(Fr *p, const uint32_t (&r)[12])
{
    uint4 *d = reinterpret_cast<uint4 *>(p);
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
}

__device__ __forceinline__ void fr_set_const(uint32_t (&r)[12], const uint32_t *c)
{
#pragma unroll
    for (int i = 0; i < 12; i++) r[i] = c[i];
}

// Montgomery -> plain integer (as_bigint, fp.tcc:270-281): multiply by 1.
__device__ __forceinline__ void fr_from_mont(uint32_t (&r)[12], const uint32_t (&a)[12])
{
    uint32_t one[12];
#pragma unroll
    for (int i = 0; i < 12; i++) one[i] = (i == 0) ? 1u : 0u;
    fr_mul_regs(r, a, one);
}

} // namespace zecale
