// Fq VM: the EC layer's execution model on sm_100a.
//
// A BW6-761 base-field element is 24 x u32. A Montgomery product is 1152 wide MADs
// (~2.4k SASS instructions, ~40 KB) and needs ~120 registers, so (a) point formulas
// cannot keep their 10+ live Fq temporaries in registers and (b) inlining a product at
// every use would make MSM kernels megabytes of SASS. Out-of-line device functions
// taking pointers to stack temporaries were tried first and miscompile under warp
// divergence (ptxas passes the uniform stack addresses in UR registers that the other
// side of a divergent branch clobbers). So every kernel here contains exactly ONE copy
// of the field arithmetic, inside a small interpreter:
//
//   * registers ("slots") are Fq elements in per-thread local memory (L1-resident,
//     hardware-interleaved so a warp's access to word k of slot s is one 128 B line);
//   * instructions are whole field operations   MUL/ADD/SUB/MOV/ONE/ZERO d, a, b
//     plus a few control ops (JZ slot, JMP, CALL/RET, a counter loop for Fermat
//     inversion) and HOOK n, which hands control to the kernel's own lambda (fetch the
//     next bucket entry, store a result, choose the next program, or halt);
//   * each thread has its own pc, so data-dependent special cases (accumulator at
//     infinity, P == Q doubling, P == -Q) are ordinary SIMT divergence inside one
//     switch — no calls, and in the common case the whole warp runs the same uop.
//
// Decode + 12 LDL.128 + 6 STL.128 per uop is < 3 % of the 1152 IMAD.WIDE of a MUL.
// The programs (point add / mixed add / double / normalise) are in fqvm_prog.cuh,
// assembled by tools/gen_fqvm_prog.py.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "field_asm.cuh"
#include "field_consts.cuh"
#include "fqvm_prog.cuh"
#include "fq_inv.cuh"

// Checked build (-DZECALE_GPU_CHECKED, `python -m zecale_b200.build --checked`): every index the sort, accumulate and
// tail kernels form and every slot number the field VM decodes is range-checked on the device; a violation prints the site and traps the kernel. Stands in for
// compute-sanitizer, which the B200 pool does not allow (profiles/r02_sanitizer.md). Compiles to nothing otherwise.
#ifdef ZECALE_GPU_CHECKED
#define ZG_CHECK(cond)                                                                                          \
    do {                                                                                                        \
        if (!(cond)) {                                                                                          \
            printf("ZG_CHECK failed: %s (%s:%d, block %u thread %u)\n", #cond, __FILE__, __LINE__, blockIdx.x, threadIdx.x); \
            __trap();                                                                                           \
        }                                                                                                       \
    } while (0)
#else
#define ZG_CHECK(cond) ((void)0)
#endif

namespace zecale {

struct alignas(16) Fq {
    uint32_t l[24];
};
struct alignas(16) Fr {
    uint32_t l[12];
};
static_assert(sizeof(Fq) == 96 && sizeof(Fr) == 48, "field element layout");

struct alignas(16) Affine {
    Fq x, y;
};
struct alignas(16) XYZZ {
    Fq x, y, zz, zzz;
};
// libff in-memory point: homogeneous projective {X, Y, Z}
struct alignas(16) Proj {
    Fq X, Y, Z;
};
static_assert(sizeof(Affine) == 192 && sizeof(XYZZ) == 384 && sizeof(Proj) == 288, "point layout");

// Slot file: VS_COUNT Fq values as uint4[6] each.
template <int N>
struct FqSlotFile {
    uint4 w[N * 6];
};
using FqSlots = FqSlotFile<VS_COUNT>;

template <class SF>
__device__ __forceinline__ void vs_load(uint32_t (&x)[24], const SF &S, uint32_t slot)
{
    ZG_CHECK(slot * 6 + 6 <= sizeof(S.w) / sizeof(uint4));
    const uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        uint4 v = p[k];
        x[4 * k] = v.x;
        x[4 * k + 1] = v.y;
        x[4 * k + 2] = v.z;
        x[4 * k + 3] = v.w;
    }
}

template <class SF>
__device__ __forceinline__ void vs_store(SF &S, uint32_t slot, const uint32_t (&x)[24])
{
    ZG_CHECK(slot * 6 + 6 <= sizeof(S.w) / sizeof(uint4));
    uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) p[k] = make_uint4(x[4 * k], x[4 * k + 1], x[4 * k + 2], x[4 * k + 3]);
}

// slot <- global/shared/any 16-byte aligned Fq
template <class SF>
__device__ __forceinline__ void vs_set(SF &S, uint32_t slot, const Fq *src)
{
    const uint4 *s = reinterpret_cast<const uint4 *>(src);
    uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) p[k] = s[k];
}

// the same for data read once (base points gathered by bucket order): ld.global.cs, evict-first, so that the stream of
// bases does not push the threads' slot files out of L2
template <class SF>
__device__ __forceinline__ void vs_set_stream(SF &S, uint32_t slot, const Fq *src)
{
    const uint4 *s = reinterpret_cast<const uint4 *>(src);
    uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) p[k] = __ldcs(s + k);
}

template <class SF>
__device__ __forceinline__ void vs_get(Fq *dst, const SF &S, uint32_t slot)
{
    uint4 *d = reinterpret_cast<uint4 *>(dst);
    const uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) d[k] = p[k];
}

template <class SF>
__device__ __forceinline__ void vs_zero(SF &S, uint32_t slot)
{
    uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) p[k] = make_uint4(0, 0, 0, 0);
}

template <class SF>
__device__ __forceinline__ void vs_one(SF &S, uint32_t slot)
{
    uint4 *p = S.w + slot * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) p[k] = make_uint4(FQ_ONE[4 * k], FQ_ONE[4 * k + 1], FQ_ONE[4 * k + 2], FQ_ONE[4 * k + 3]);
}

template <class SF>
__device__ __forceinline__ bool vs_is_zero(const SF &S, uint32_t slot)
{
    const uint4 *p = S.w + slot * 6;
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        uint4 v = p[k];
        acc |= v.x | v.y | v.z | v.w;
    }
    return acc == 0;
}

// slot <- -(*src) mod q (0 stays 0); cf. fp.tcc:617-630
template <bool STREAM = false, class SF>
__device__ __forceinline__ void vs_set_neg(SF &S, uint32_t slot, const Fq *src)
{
    uint32_t a[24], z[24], r[24];
    const uint4 *s = reinterpret_cast<const uint4 *>(src);
#pragma unroll
    for (int k = 0; k < 6; k++) {
        uint4 v = STREAM ? __ldcs(s + k) : s[k];
        a[4 * k] = v.x;
        a[4 * k + 1] = v.y;
        a[4 * k + 2] = v.z;
        a[4 * k + 3] = v.w;
        z[4 * k] = z[4 * k + 1] = z[4 * k + 2] = z[4 * k + 3] = 0;
    }
    fq_sub_regs(r, z, a);
    vs_store(S, slot, r);
}

// Accumulator (slots X..ZZZ) <-> memory
template <class SF>
__device__ __forceinline__ void vs_store_acc(XYZZ *dst, const SF &S)
{
    vs_get(&dst->x, S, VS_X);
    vs_get(&dst->y, S, VS_Y);
    vs_get(&dst->zz, S, VS_ZZ);
    vs_get(&dst->zzz, S, VS_ZZZ);
}
template <class SF>
__device__ __forceinline__ void vs_load_q(SF &S, const XYZZ *src)
{
    vs_set(S, VS_QX, &src->x);
    vs_set(S, VS_QY, &src->y);
    vs_set(S, VS_QZZ, &src->zz);
    vs_set(S, VS_QZZZ, &src->zzz);
}
template <class SF>
__device__ __forceinline__ void vs_load_acc(SF &S, const XYZZ *src)
{
    vs_set(S, VS_X, &src->x);
    vs_set(S, VS_Y, &src->y);
    vs_set(S, VS_ZZ, &src->zz);
    vs_set(S, VS_ZZZ, &src->zzz);
}
template <class SF>
__device__ __forceinline__ void vs_acc_inf(SF &S)
{
    vs_zero(S, VS_X);
    vs_one(S, VS_Y);
    vs_zero(S, VS_ZZ);
    vs_zero(S, VS_ZZZ);
}
__device__ __forceinline__ void xyzz_store_inf(XYZZ *dst)
{
    uint4 *d = reinterpret_cast<uint4 *>(dst);
#pragma unroll
    for (int k = 0; k < 24; k++) d[k] = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int k = 0; k < 6; k++) d[6 + k] = make_uint4(FQ_ONE[4 * k], FQ_ONE[4 * k + 1], FQ_ONE[4 * k + 2], FQ_ONE[4 * k + 3]);
}

// ---------------------------------------------------------------- plain modular inverse
// r = a^-1 mod q for a plain integer 0 < a < q: binary extended Euclid (shift / subtract).
// Invariants: u = x1*a, v = x2*a (mod q); ends when u or v reaches 1. ~1500 iterations of
// 24-limb adds/shifts: ~60 Montgomery-product times, against 1141 products for a^(q-2).
// a == 0 returns 0.
__device__ __forceinline__ void big_shr1(uint32_t (&x)[24], uint32_t top)
{
#pragma unroll
    for (int k = 0; k < 23; k++) x[k] = __funnelshift_r(x[k], x[k + 1], 1);
    x[23] = __funnelshift_r(x[23], top, 1);
}

// x = x/2 mod q
__device__ __forceinline__ void big_halve_mod(uint32_t (&x)[24])
{
    uint32_t carry = 0;
    if (x[0] & 1u) {
        uint64_t c = 0;
#pragma unroll
        for (int k = 0; k < 24; k++) {
            c += (uint64_t)x[k] + FQ_P[k];
            x[k] = (uint32_t)c;
            c >>= 32;
        }
        carry = (uint32_t)c;
    }
    big_shr1(x, carry);
}

// d = x - y, returns the borrow
__device__ __forceinline__ uint32_t big_sub(uint32_t (&d)[24], const uint32_t (&x)[24], const uint32_t (&y)[24])
{
    int64_t c = 0;
#pragma unroll
    for (int k = 0; k < 24; k++) {
        c += (int64_t)x[k] - (int64_t)y[k];
        d[k] = (uint32_t)c;
        c >>= 32;
    }
    return (uint32_t)(c & 1);
}

__device__ __forceinline__ bool big_is_one(const uint32_t (&x)[24])
{
    uint32_t acc = x[0] ^ 1u;
#pragma unroll
    for (int k = 1; k < 24; k++) acc |= x[k];
    return acc == 0;
}

__device__ __forceinline__ void fq_inv_plain(uint32_t (&r)[24], const uint32_t (&a)[24])
{
    uint32_t u[24], v[24], x1[24], x2[24], d[24];
    uint32_t nz = 0;
#pragma unroll
    for (int k = 0; k < 24; k++) {
        u[k] = a[k];
        v[k] = FQ_P[k];
        x1[k] = k == 0 ? 1u : 0u;
        x2[k] = 0;
        nz |= a[k];
    }
    if (nz == 0) {
#pragma unroll
        for (int k = 0; k < 24; k++) r[k] = 0;
        return;
    }
    for (;;) {
        const bool u1 = big_is_one(u);
        if (u1 || big_is_one(v)) {
#pragma unroll
            for (int k = 0; k < 24; k++) r[k] = u1 ? x1[k] : x2[k];
            return;
        }
        if (!(u[0] & 1u)) {
            big_shr1(u, 0);
            big_halve_mod(x1);
        } else if (!(v[0] & 1u)) {
            big_shr1(v, 0);
            big_halve_mod(x2);
        } else if (big_sub(d, u, v) == 0) { // u >= v
#pragma unroll
            for (int k = 0; k < 24; k++) u[k] = d[k];
            fq_sub_regs(x1, x1, x2);
        } else {
            big_sub(v, v, u);
            fq_sub_regs(x2, x2, x1);
        }
    }
}

// Run from `pc` until the hook returns false. hook(id, pc, S) may change pc.
// Slot operands with bit 7 set are relative: slot = (operand & 0x7f) + rel (the hook moves `rel`
// to walk an array of per-chain slot groups with one program).
// `wide` (optional): where MULW parks the unreduced 1536-bit product until MULFIN, as 12 uint4 `wide_stride` elements apart
// - shared memory in k_accumulate (one column per thread); nullptr = the W0/W1 slot pair of the thread's own slot file.
template <bool REL, bool INV, class SF, class Hook>
__device__ __forceinline__ void fqvm_run_impl(SF &S, uint32_t pc, Hook &&hook, const uint32_t &rel, uint4 *wide = nullptr,
                                              uint32_t wide_stride = 1)
{
    uint32_t ret = 0, cnt = 0;
    auto rs = [&](uint32_t s) -> uint32_t { return (REL && (s & 0x80u)) ? (s & 0x7fu) + rel : s; };
    for (;;) {
        const uint32_t u = FQVM_CODE[pc++];
        const uint32_t op = u & 0xffu, imm = u >> 16;
        const uint32_t d = op <= VOP_JZ ? rs((u >> 8) & 0xffu) : (u >> 8) & 0xffu;
        if (op <= VOP_MULFIN) {
            uint32_t x[24], y[24], z[24];
            vs_load(x, S, rs(imm & 0xffu));
            vs_load(y, S, rs(imm >> 8));
            if (op == VOP_ADD)
                fq_add_regs(z, x, y);
            else if (op == VOP_SUB)
                fq_sub_regs(z, x, y);
            else {
                // product and reduction are separate bodies (field_asm.cuh), one copy of each per kernel: MUL = mulw + redc;
                // MULW c, e stashes the unreduced c*e in the wide slot pair; MULFIN d, a, b reduces a*b - c*e + q^2 ONCE.
                uint32_t w[48];
                fq_mulw_regs(w, x, y);
                uint4 *ws = wide ? wide : S.w + VS_W0 * 6;
                const uint32_t wst = wide ? wide_stride : 1u;
                if (op == VOP_MULW) {
#pragma unroll
                    for (int k = 0; k < 12; k++) ws[k * wst] = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
                    continue;
                }
                if (op == VOP_MULFIN) {
                    uint32_t t[48];
#pragma unroll
                    for (int k = 0; k < 12; k++) {
                        const uint4 v = ws[k * wst];
                        t[4 * k] = v.x;
                        t[4 * k + 1] = v.y;
                        t[4 * k + 2] = v.z;
                        t[4 * k + 3] = v.w;
                    }
                    fq_wsubk_regs(w, t);
                }
                fq_redc_regs(z, w);
            }
            vs_store(S, d, z);
            continue;
        }
        switch (op) {
        case VOP_MOV: {
            const uint4 *s = S.w + rs(imm) * 6;
            uint4 *p = S.w + d * 6;
#pragma unroll
            for (int k = 0; k < 6; k++) p[k] = s[k];
            break;
        }
        case VOP_ONE: vs_one(S, d); break;
        case VOP_ZERO: vs_zero(S, d); break;
        case VOP_JZ:
            if (vs_is_zero(S, d)) pc = imm;
            break;
        case VOP_JMP: pc = imm; break;
        case VOP_JBIT:
            if ((FQ_PM2[cnt >> 5] >> (cnt & 31u)) & 1u) pc = imm;
            break;
        case VOP_DECJNZ:
            if (cnt != 0) {
                cnt--;
                pc = imm;
            }
            break;
        case VOP_SETCNT: cnt = imm; break;
        case VOP_CALL:
            ret = pc;
            pc = imm;
            break;
        case VOP_RET: pc = ret; break;
        case VOP_INVP: {
            if (!INV) return; // kernels whose programs never invert leave the inversion code out
            uint32_t x[24], z[24];
            vs_load(x, S, imm);
            fq_inv_bingcd(z, x, FQ_P, FQ_INV32);
            vs_store(S, d, z);
            break;
        }
        case VOP_LDR3: {
            uint4 *p = S.w + d * 6;
#pragma unroll
            for (int k = 0; k < 6; k++) p[k] = make_uint4(FQ_R3[4 * k], FQ_R3[4 * k + 1], FQ_R3[4 * k + 2], FQ_R3[4 * k + 3]);
            break;
        }
        case VOP_HOOK:
            if (!hook(d, pc, S)) return;
            break;
        default: return; // HALT
        }
    }
}

template <class SF, class Hook>
__device__ __forceinline__ void fqvm_run(SF &S, uint32_t pc, Hook &&hook, const uint32_t &rel)
{
    fqvm_run_impl<true, true>(S, pc, hook, rel);
}
template <class SF, class Hook>
__device__ __forceinline__ void fqvm_run(SF &S, uint32_t pc, Hook &&hook)
{
    const uint32_t rel0 = 0;
    fqvm_run_impl<false, true>(S, pc, hook, rel0);
}
// for kernels that only run point additions / doublings (no INVP in their programs)
template <class SF, class Hook>
__device__ __forceinline__ void fqvm_run_noinv(SF &S, uint32_t pc, Hook &&hook, uint4 *wide = nullptr, uint32_t wide_stride = 1)
{
    const uint32_t rel0 = 0;
    fqvm_run_impl<false, false>(S, pc, hook, rel0, wide, wide_stride);
}

// ---------------------------------------------------------------- Fr (register form)
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
