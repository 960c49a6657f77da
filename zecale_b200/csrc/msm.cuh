// Pippenger bucket MSM over BW6-761 G1/G2 for sm_100a.
//
// Replaces libff::multi_exp<G, Fr, multi_exp_method_BDLO12_signed, special>
// (libff/algebra/scalar_multiplication/multiexp.tcc:507-688) and its callers
// multi_exp_filter_one_zero (:695-757) / kc_multi_exp_with_mixed_addition
// (libsnark knowledge_commitment/kc_multiexp.tcc:31-113).
//
// The reference walks one c-bit window at a time, top-down, on one OpenMP chunk per
// thread, with c doublings between windows. Here:
//
//   * the W signed c-bit windows of all scalars are decomposed at once (k_digits;
//     field_get_signed_digit semantics, field_utils.tcc:81-115) and counting-sorted by
//     bucket (histogram atomics -> scan -> k_scatter);
//   * window w = g*V + v is served by base copy g (the pinned image holds
//     2^(c*V*g) * P_i for g < G, computed once at pin time) and bucket set v, so only V
//     bucket sets exist and the final Horner needs c*(V-1) doublings — none when the
//     proving key is fully expanded (V = 1). An un-pinned call uses G = 1, V = W;
//   * buckets are cut into tasks of <= S entries, tasks are sorted by length so the 32
//     lanes of a warp run equally long chains of mixed additions (k_accumulate, one
//     thread per task, XYZZ accumulator in the Fq VM);
//   * sum_b (b+1) B_b (multiexp_accumulate_buckets, :90-125) is a binary tree over the bucket
//     index with one bit tree per index bit riding along (k_tail_level: 2 NB additions, depth
//     c-1), then parallel doublings, a pair-tree sum and the normalisation to libff special
//     form; every launch of the tail serves all MSMs of a proof (TailBatch).
//
// No 0/1 filter is needed: a zero scalar has all-zero digits, a one is an ordinary
// bucket entry, and base points at infinity never produce entries.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "fqvm.cuh"

namespace zecale {

// resident CTAs per SM of k_accumulate (register cap 65536 / (128 * ZG_ACC_CTAS))
#ifndef ZG_ACC_CTAS
#define ZG_ACC_CTAS 3
#endif
// k_accumulate parks the unreduced product of MULW / MULFIN in shared memory (24 KB per CTA) instead of the thread's
// local-memory slot file: same speed, DRAM writes of a launch 1.00 -> 0.40 GB (profiles/r02_traffic.json). 0 = slot file.
#ifndef ZG_ACC_SHARED_STASH
#define ZG_ACC_SHARED_STASH 1
#endif

static constexpr int FR_SCALAR_BITS = 377;
static constexpr uint32_t MSM_MAX_TASKS_PER_BUCKET = 256;
static constexpr uint64_t MSM_TARGET_TASKS = 3ull * 148 * 3 * 128;

struct MsmPlan {
    uint32_t n;            // (base, scalar) pairs
    uint32_t c;            // window bits
    uint32_t W;            // windows = ceil(378 / c)
    uint32_t V;            // bucket sets
    uint32_t G;            // base copies used = ceil(W / V)
    uint32_t NB;           // buckets per set = 2^(c-1)
    uint32_t S;            // max entries per accumulate task
    uint32_t copy_stride;  // points per base copy
    uint64_t n_buckets;    // V * NB
    uint64_t max_tasks;
    uint32_t acc_mode;     // MSM_ACC_XYZZ | MSM_ACC_TREE
    uint32_t tree_rounds;  // pair-tree rounds launched = ceil(log2(n * G)) (a bucket holds <= n*G entries)
    uint32_t tree_min;     // rounds every bucket goes through before the XYZZ chains take over (k_tree_chain)
    uint64_t n_entries;    // n * W: capacity of digits / entries
    uint64_t n_bases;      // G * copy_stride: points of the (expanded) base image
};

// Bucket accumulation mode: ZECALE_GPU_ACC = "tree" (batched affine pair tree, k_tree_round) or
// "xyzz" (one XYZZ chain per task, k_accumulate). Read at every plan so tests can switch it.
static constexpr uint32_t MSM_ACC_XYZZ = 0, MSM_ACC_TREE = 1;
static constexpr uint32_t MSM_ACC_DEFAULT = MSM_ACC_XYZZ;
inline uint32_t msm_acc_mode()
{
    const char *v = getenv("ZECALE_GPU_ACC");
    if (v && v[0] == 'x') return MSM_ACC_XYZZ;
    if (v && v[0] == 't') return MSM_ACC_TREE;
    return MSM_ACC_DEFAULT;
}
// Hybrid accumulation: the first rounds of the pair tree hold almost all additions (1/2, 1/4, 1/8 ...
// of them) in big, perfectly balanced batches of affine additions (6 products each + a shared
// inversion); what is left after MSM_TREE_MIN_ROUNDS rounds - <= ceil(n_b / 2^R) points per bucket -
// is summed by one XYZZ chain per bucket (10 products per addition, no batch to fill). Buckets much
// heavier than the rest keep halving until no chain is longer than MSM_TREE_CHAIN_MAX points.
static constexpr uint32_t MSM_TREE_MIN_ROUNDS = 3;
static constexpr uint32_t MSM_TREE_CHAIN_MAX = 16;
inline uint32_t msm_tree_min_rounds()
{
    const char *v = getenv("ZECALE_GPU_TREE_ROUNDS");
    return v ? (uint32_t)atoi(v) : MSM_TREE_MIN_ROUNDS;
}

struct TreeMeta {
    uint32_t max_count; // largest bucket (set by k_tree_count)
};

// Window sizes the bucket reduction (k_tail_*) is written for.
static constexpr uint32_t MSM_MIN_C = 3, MSM_MAX_C = 20;

inline uint32_t msm_default_c(uint32_t n)
{
    uint32_t lg = 0;
    while ((1ull << (lg + 1)) <= n) lg++;
    int c = (int)lg - 3;
    if (c < 4) c = 4;
    if (c > 16) c = 16;
    return (uint32_t)c;
}

// c, V fixed by the caller (pinned image) or defaulted (V = W).
inline MsmPlan msm_make_plan(uint32_t n, uint32_t c, uint32_t V, uint32_t copy_stride)
{
    MsmPlan p{};
    p.n = n;
    p.c = c;
    p.W = (FR_SCALAR_BITS + 1 + c - 1) / c;
    p.V = V ? (V > p.W ? p.W : V) : p.W;
    p.G = (p.W + p.V - 1) / p.V;
    p.NB = 1u << (c - 1);
    p.copy_stride = copy_stride;
    // Task length S (max entries per accumulate task): long enough that an ordinary bucket
    // (mean + 4 sigma of a Poisson load) is ONE task — a mixed addition inside a task costs 10
    // products, folding two partial sums 14 — and at least what keeps the task count near 3 waves
    // of the 148 SM x 3 CTA x 128 thread accumulate grid for very large n. SM load balance comes
    // from the length-sorted task order and from the sibling MSMs of a proof running concurrently.
    // Buckets that do get cut (a Poisson tail, many equal small scalars, or the few buckets a short
    // top window feeds) are folded by k_tail_fold, one warp per bucket.
    uint64_t avg = ((uint64_t)n * p.W + (uint64_t)p.V * p.NB - 1) / ((uint64_t)p.V * p.NB);
    uint64_t S = ((uint64_t)n * p.W + MSM_TARGET_TASKS - 1) / MSM_TARGET_TASKS;
    uint64_t sq = 0;
    while ((sq + 1) * (sq + 1) <= avg) sq++;
    if (S < avg + 4 * sq + 8) S = avg + 4 * sq + 8;
    p.S = (uint32_t)S;
    p.acc_mode = msm_acc_mode();
    p.tree_rounds = 0;
    while ((1ull << p.tree_rounds) < (uint64_t)n * p.G) p.tree_rounds++;
    p.tree_min = msm_tree_min_rounds();
    p.n_buckets = (uint64_t)p.V * p.NB;
    p.max_tasks = ((uint64_t)n * p.W) / p.S + p.n_buckets + 1;
    p.n_entries = (uint64_t)n * p.W;
    p.n_bases = (uint64_t)p.G * copy_stride;
    return p;
}

// ------------------------------------------------------------------ import
// libff special-form points (stride bytes apart, {X,Y,Z}) -> device affine image + flags.
// status bit0 is raised if a base is neither Z == one nor Z == 0.
__global__ void k_import_special(const uint8_t *raw, size_t stride, uint32_t n, Affine *out, uint8_t *is_inf, int *status)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 *src = reinterpret_cast<const uint4 *>(raw + (size_t)i * stride);
    uint4 z[6];
    uint32_t zor = 0, zne = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        z[k] = src[12 + k];
        zor |= z[k].x | z[k].y | z[k].z | z[k].w;
        zne |= (z[k].x ^ FQ_ONE[4 * k]) | (z[k].y ^ FQ_ONE[4 * k + 1]) | (z[k].z ^ FQ_ONE[4 * k + 2]) | (z[k].w ^ FQ_ONE[4 * k + 3]);
    }
    uint4 *dst = reinterpret_cast<uint4 *>(out + i);
    if (zor == 0) {
#pragma unroll
        for (int k = 0; k < 12; k++) dst[k] = make_uint4(0, 0, 0, 0);
        is_inf[i] = 1;
        return;
    }
    if (zne != 0) atomicOr(status, 1);
    is_inf[i] = 0;
#pragma unroll
    for (int k = 0; k < 12; k++) dst[k] = src[k];
}

// Key-file points -> device affine image + flags. libzeth writes a group element as
// X || Y, each coordinate the PLAIN integer in big-endian bytes (group_element_write_bytes,
// libzeth/core/group_element_utils.tcc:93-100 -> libff field_write<encoding_binary, form_plain>,
// field_serialization.tcc:197-207); zero is (0, 1) (group_element_read_bytes :102-113). `raw` is
// 4-byte aligned, points `stride` bytes apart. Byte-reverse, then one Montgomery product with R^2 per
// coordinate — what Fp_model(bigint) does on the host (fp.tcc:246-252).
__global__ void __launch_bounds__(128) k_import_bytes(const uint8_t *raw, size_t stride, uint32_t n, Affine *out, uint8_t *is_inf)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(raw + (size_t)i * stride);
    uint32_t x[24], y[24], r2[24], mx[24], my[24];
    uint32_t xor_ = 0, yne = 0;
#pragma unroll
    for (int k = 0; k < 24; k++) {
        x[k] = __byte_perm(src[23 - k], 0, 0x0123);
        y[k] = __byte_perm(src[47 - k], 0, 0x0123);
        r2[k] = FQ_R2[k];
        xor_ |= x[k];
        yne |= y[k] ^ (k == 0 ? 1u : 0u);
    }
    uint4 *dst = reinterpret_cast<uint4 *>(out + i);
    if (xor_ == 0 && yne == 0) {
#pragma unroll
        for (int k = 0; k < 12; k++) dst[k] = make_uint4(0, 0, 0, 0);
        is_inf[i] = 1;
        return;
    }
    is_inf[i] = 0;
    fq_mul_regs(mx, x, r2);
    fq_mul_regs(my, y, r2);
#pragma unroll
    for (int k = 0; k < 6; k++) {
        dst[k] = make_uint4(mx[4 * k], mx[4 * k + 1], mx[4 * k + 2], mx[4 * k + 3]);
        dst[6 + k] = make_uint4(my[4 * k], my[4 * k + 1], my[4 * k + 2], my[4 * k + 3]);
    }
}

// ------------------------------------------------------------------ synthetic bases
// out[i] = k_i * generator in libff special form, k_i = 32-bit hash of (seed, i), never 0.
// Used by bench.py / tests to make distinct valid curve points without a CPU library.
__device__ __forceinline__ uint32_t synth_scalar(uint64_t seed, uint32_t i)
{
    uint64_t x = seed * 0x9E3779B97F4A7C15ull + (uint64_t)i * 0xBF58476D1CE4E5B9ull + 0x94D049BB133111EBull;
    x ^= x >> 30;
    x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27;
    x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    uint32_t k = (uint32_t)x;
    return k ? k : 1u;
}

__global__ void __launch_bounds__(128, 2) k_synth_bases(int group, uint64_t seed, uint32_t n, Proj *out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    FqSlots S;
    vs_acc_inf(S);
    {
        const uint32_t *gx = group == 2 ? G2_GEN_X : G1_GEN_X, *gy = group == 2 ? G2_GEN_Y : G1_GEN_Y;
        uint32_t x[24], y[24];
#pragma unroll
        for (int k = 0; k < 24; k++) {
            x[k] = gx[k];
            y[k] = gy[k];
        }
        vs_store(S, VS_QX, x);
        vs_store(S, VS_QY, y);
    }
    const uint32_t k = synth_scalar(seed, i);
    int bit = 31;
    uint32_t phase = 0; // 0: a doubling just ran, 1: an addition just ran, 2: normalising
    fqvm_run(S, VPC_DBL, [&](uint32_t, uint32_t &pc, FqSlots &S_) {
        if (phase == 2) {
            Proj *o = out + i;
            vs_get(&o->X, S_, VS_QX);
            vs_get(&o->Y, S_, VS_QY);
            vs_one(S_, VS_T0);
            vs_get(&o->Z, S_, VS_T0);
            return false;
        }
        if (phase == 0 && ((k >> bit) & 1u)) {
            phase = 1;
            pc = VPC_MADD;
            return true;
        }
        phase = 0;
        if (bit == 0) {
            phase = 2;
            pc = VPC_TOAFF; // k != 0 and the generator has prime order: never infinity
            return true;
        }
        bit--;
        pc = VPC_DBL;
        return true;
    });
}

// ------------------------------------------------------------------ pinned-image expansion
// dst[i] = 2^ndbl * src[i] (affine in, affine out; (0,0) = infinity stays infinity). Each thread
// owns K consecutive points and shares one inversion among them (Montgomery's trick):
//   forward : XYZZ_j = 2^ndbl * P_j -> scratch, prefix_j = prod_{j' < j} ZZZ_j' -> scratch
//   invert  : 1 / prod_j ZZZ_j            (one binary-Euclid inversion per K points)
//   backward: 1/ZZZ_j = inv * prefix_j ; inv *= ZZZ_j ; x = X/ZZ, y = Y/ZZZ
// This is libff's batch_to_special (multiexp.tcc:949-974, batch_invert field_utils.tcc:330-350)
// fused with the doublings.
static constexpr uint32_t EXPAND_K = 8;

__global__ void __launch_bounds__(128, 2)
k_expand_copy(const Affine *src, Affine *dst, uint32_t n, uint32_t ndbl, XYZZ *scratch_pt, Fq *scratch_prefix)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i0 = t * EXPAND_K;
    if (i0 >= n) return;
    const uint32_t i1 = i0 + EXPAND_K < n ? i0 + EXPAND_K : n;
    FqSlots S;
    uint32_t i = i0, left = 0, stage = 0; // stage 0 forward, 1 inverting, 2 backward
    vs_one(S, VS_K1);
    // forward: find the next finite point, load it, start its doublings
    auto next_forward = [&](uint32_t &pc, FqSlots &S_) -> bool {
        while (i < i1) {
            const Affine *p = src + i;
            vs_set(S_, VS_X, &p->x);
            vs_set(S_, VS_Y, &p->y);
            if (!(vs_is_zero(S_, VS_X) && vs_is_zero(S_, VS_Y))) {
                vs_one(S_, VS_ZZ);
                vs_one(S_, VS_ZZZ);
                left = ndbl;
                pc = VPC_DBL;
                return true;
            }
            xyzz_store_inf(scratch_pt + i); // zz = 0 marks infinity
            i++;
        }
        return false;
    };
    auto next_backward = [&](uint32_t &pc, FqSlots &S_) -> bool {
        while (i > i0) {
            i--;
            vs_load_acc(S_, scratch_pt + i);
            if (!vs_is_zero(S_, VS_ZZ)) {
                vs_set(S_, VS_T3, scratch_prefix + i);
                pc = VPC_UNSTEP;
                return true;
            }
            uint4 *d = reinterpret_cast<uint4 *>(dst + i);
            for (int k = 0; k < 12; k++) d[k] = make_uint4(0, 0, 0, 0);
        }
        return false;
    };
    uint32_t pc0 = 0;
    if (!next_forward(pc0, S)) { // all K points at infinity
        stage = 2;
        i = i1;
        next_backward(pc0, S);
        return;
    }
    fqvm_run(S, pc0, [&](uint32_t, uint32_t &pc, FqSlots &S_) {
        if (stage == 0) {
            if (left) { // a doubling finished
                if (--left) {
                    pc = VPC_DBL;
                    return true;
                }
                if (vs_is_zero(S_, VS_ZZZ)) { // doubled to infinity (2-torsion): keep it out of the product
                    xyzz_store_inf(scratch_pt + i);
                    i++;
                } else {
                    vs_store_acc(scratch_pt + i, S_);
                    vs_get(scratch_prefix + i, S_, VS_K1);
                    pc = VPC_PRODSTEP; // T2 *= ZZZ
                    return true;
                }
            } else {
                i++; // prodstep finished
            }
            if (next_forward(pc, S_)) return true;
            stage = 1;
            pc = VPC_INVPROD; // T0 = 1 / T2
            return true;
        }
        if (stage == 1) {
            stage = 2;
            i = i1;
            return next_backward(pc, S_);
        }
        // an unstep finished: affine point in QX, QY
        vs_get(&dst[i].x, S_, VS_QX);
        vs_get(&dst[i].y, S_, VS_QY);
        return next_backward(pc, S_);
    });
}

// ------------------------------------------------------------------ digits
// One thread per scalar. scalars are Montgomery Fr, `sstride` bytes apart; the scalar of
// base i is scalars[index ? index[i] : i]. digits[w*n + i] = signed digit of window w.
__global__ void k_digits(const uint8_t *scalars, size_t sstride, const uint64_t *index, const uint8_t *is_inf, MsmPlan p,
                         int32_t *digits, uint32_t *counts)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    size_t si = index ? (size_t)index[i] : (size_t)i;
    uint32_t m[12];
    __align__(8) uint32_t v[14];
    fr_load(m, reinterpret_cast<const Fr *>(scalars + si * sstride));
    {
        uint32_t t[12];
        fr_from_mont(t, m);
#pragma unroll
        for (int k = 0; k < 12; k++) v[k] = t[k];
        v[12] = 0;
        v[13] = 0;
    }
    const bool skip = is_inf[i] != 0;
    const uint32_t c = p.c;
    const uint32_t mask = (1u << c) - 1;
    const uint32_t half = 1u << (c - 1);
    uint32_t carry = 0;
    for (uint32_t w = 0; w < p.W; w++) {
        uint32_t bit = w * c;
        uint32_t lo = bit >> 5, sh = bit & 31;
        ZG_CHECK(lo + 1 < 14);
        uint64_t two = (uint64_t)v[lo] | ((uint64_t)v[lo + 1] << 32);
        uint32_t d = (uint32_t)(two >> sh) & mask;
        d += carry;
        int32_t sd;
        if (d > half) {
            sd = (int32_t)d - (int32_t)(1u << c);
            carry = 1;
        } else {
            sd = (int32_t)d;
            carry = 0;
        }
        if (skip) sd = 0;
        digits[(size_t)w * p.n + i] = sd;
        if (sd != 0) {
            uint32_t b = (sd < 0 ? (uint32_t)(-sd) : (uint32_t)sd) - 1;
            ZG_CHECK(b < p.NB);
            atomicAdd(&counts[(size_t)(w % p.V) * p.NB + b], 1u);
        }
    }
}

__global__ void k_scatter(MsmPlan p, const int32_t *digits, uint32_t *cursor, uint32_t *entries)
{
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t total = (size_t)p.n * p.W;
    if (t >= total) return;
    uint32_t w = (uint32_t)(t / p.n);
    uint32_t i = (uint32_t)(t - (size_t)w * p.n);
    int32_t sd = digits[t];
    if (sd == 0) return;
    uint32_t neg = sd < 0;
    uint32_t b = (neg ? (uint32_t)(-sd) : (uint32_t)sd) - 1;
    ZG_CHECK(b < p.NB);
    uint32_t pos = atomicAdd(&cursor[(size_t)(w % p.V) * p.NB + b], 1u);
    ZG_CHECK(pos < p.n_entries && (uint64_t)(w / p.V) * p.copy_stride + i < p.n_bases);
    entries[pos] = ((w / p.V) * p.copy_stride + i) | (neg << 31);
}

// ------------------------------------------------------------------ scan (uint32, exclusive)
static constexpr int SCAN_BLOCK = 1024;
static constexpr int SCAN_ITEMS = 4; // per thread

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *smem, uint32_t &total)
{
    int tid = threadIdx.x;
    smem[tid] = v;
    __syncthreads();
    for (int d = 1; d < blockDim.x; d <<= 1) {
        uint32_t t = tid >= d ? smem[tid - d] : 0;
        __syncthreads();
        smem[tid] += t;
        __syncthreads();
    }
    total = smem[blockDim.x - 1];
    uint32_t incl = smem[tid];
    __syncthreads();
    return incl - v;
}

__global__ void k_scan_reduce(const uint32_t *in, size_t n, uint32_t *block_sums)
{
    __shared__ uint32_t sm[SCAN_BLOCK];
    size_t base = (size_t)blockIdx.x * SCAN_BLOCK * SCAN_ITEMS;
    uint32_t s = 0;
    for (int k = 0; k < SCAN_ITEMS; k++) {
        size_t idx = base + (size_t)threadIdx.x * SCAN_ITEMS + k;
        if (idx < n) s += in[idx];
    }
    uint32_t total;
    block_exclusive_scan(s, sm, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

__global__ void k_scan_sums(uint32_t *block_sums, size_t nb, uint32_t *grand_total)
{
    __shared__ uint32_t sm[SCAN_BLOCK];
    uint32_t running = 0;
    for (size_t base = 0; base < nb; base += SCAN_BLOCK) {
        size_t idx = base + threadIdx.x;
        uint32_t v = idx < nb ? block_sums[idx] : 0;
        uint32_t total;
        uint32_t ex = block_exclusive_scan(v, sm, total);
        if (idx < nb) block_sums[idx] = running + ex;
        running += total;
    }
    if (threadIdx.x == 0 && grand_total) *grand_total = running;
}

__global__ void k_scan_apply(const uint32_t *in, size_t n, const uint32_t *block_sums, uint32_t *out)
{
    __shared__ uint32_t sm[SCAN_BLOCK];
    size_t base = (size_t)blockIdx.x * SCAN_BLOCK * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
    for (int k = 0; k < SCAN_ITEMS; k++) {
        size_t idx = base + (size_t)threadIdx.x * SCAN_ITEMS + k;
        v[k] = idx < n ? in[idx] : 0;
        s += v[k];
    }
    uint32_t total;
    uint32_t ex = block_exclusive_scan(s, sm, total) + block_sums[blockIdx.x];
    for (int k = 0; k < SCAN_ITEMS; k++) {
        size_t idx = base + (size_t)threadIdx.x * SCAN_ITEMS + k;
        if (idx < n) out[idx] = ex;
        ex += v[k];
    }
}

// out[0..n) = exclusive scan of in[0..n); out[n] = total. `tmp` holds ceil(n/4096)+1 words.
inline void exclusive_scan_u32(const uint32_t *in, size_t n, uint32_t *out, uint32_t *tmp, cudaStream_t st)
{
    size_t nb = (n + (size_t)SCAN_BLOCK * SCAN_ITEMS - 1) / ((size_t)SCAN_BLOCK * SCAN_ITEMS);
    k_scan_reduce<<<(unsigned)nb, SCAN_BLOCK, 0, st>>>(in, n, tmp);
    k_scan_sums<<<1, SCAN_BLOCK, 0, st>>>(tmp, nb, out + n);
    k_scan_apply<<<(unsigned)nb, SCAN_BLOCK, 0, st>>>(in, n, tmp, out);
}

// ------------------------------------------------------------------ tasks
__device__ __forceinline__ uint32_t msm_bucket_tasks(uint32_t cnt, uint32_t S)
{
    uint32_t nt = (cnt + S - 1) / S;
    return nt > MSM_MAX_TASKS_PER_BUCKET ? MSM_MAX_TASKS_PER_BUCKET : nt;
}

__global__ void k_task_count(MsmPlan p, const uint32_t *counts, uint32_t *ntasks)
{
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.n_buckets) return;
    ntasks[b] = msm_bucket_tasks(counts[b], p.S);
}

// Pair-tree mode: one "task" (= one partial sum) per non-empty bucket, and the largest bucket size.
__global__ void k_tree_count(MsmPlan p, const uint32_t *counts, uint32_t *ntasks, TreeMeta *meta)
{
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t c = b < p.n_buckets ? counts[b] : 0;
    if (b < p.n_buckets) ntasks[b] = c ? 1u : 0u;
    // one atomic per warp
    for (int d = 16; d; d >>= 1) {
        uint32_t o = __shfl_xor_sync(0xffffffffu, c, d);
        c = o > c ? o : c;
    }
    if ((threadIdx.x & 31) == 0 && c) atomicMax(&meta->max_count, c);
}

// Task t of bucket b covers entries [start, start+len). size_hist bins tasks by length
// (capped) for the length sort.
static constexpr uint32_t MSM_LEN_BINS = 1024;
__device__ __forceinline__ uint32_t msm_len_bin(uint32_t len) { return len >= MSM_LEN_BINS ? 0u : MSM_LEN_BINS - len; }

__global__ void k_task_fill(MsmPlan p, const uint32_t *counts, const uint32_t *offsets, const uint32_t *task_off,
                            uint32_t *task_start, uint32_t *task_len, uint32_t *task_bucket, uint32_t *len_hist,
                            uint32_t *heavy_list, uint32_t *heavy_count)
{
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.n_buckets) return;
    uint32_t t0 = task_off[b], nt = task_off[b + 1] - t0;
    if (nt > 1) heavy_list[atomicAdd(heavy_count, 1u)] = (uint32_t)b; // folded by k_tail_fold
    uint64_t cnt = counts[b];
    uint32_t off = offsets[b];
    ZG_CHECK((uint64_t)t0 + nt <= p.max_tasks && off + cnt <= p.n_entries && (nt > 0) == (cnt > 0));
    for (uint32_t k = 0; k < nt; k++) {
        uint32_t lo = (uint32_t)((cnt * k) / nt), hi = (uint32_t)((cnt * (k + 1)) / nt);
        task_start[t0 + k] = off + lo;
        task_len[t0 + k] = hi - lo;
        task_bucket[t0 + k] = (uint32_t)b;
        atomicAdd(&len_hist[msm_len_bin(hi - lo)], 1u);
    }
}

// order[pos] = task id, longest tasks first.
__global__ void k_task_order(const uint32_t *task_off_total, const uint32_t *task_len, uint32_t *len_cursor, uint32_t *order)
{
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *task_off_total) return;
    uint32_t pos = atomicAdd(&len_cursor[msm_len_bin(task_len[t])], 1u);
    ZG_CHECK(pos < *task_off_total);
    order[pos] = t;
}

// ------------------------------------------------------------------ accumulate
// One thread per task: partial[t] = sum of its entries (mixed additions).
__global__ void __launch_bounds__(128, ZG_ACC_CTAS)
k_accumulate(MsmPlan p, const Affine *bases, const uint32_t *n_tasks, const uint32_t *order, const uint32_t *task_start,
             const uint32_t *task_len, const uint32_t *entries, XYZZ *partial)
{
#if ZG_ACC_SHARED_STASH
    __shared__ uint4 stash[12 * 128]; // the wide product of MULW / MULFIN: column threadIdx.x, conflict-free 16-byte accesses
#endif
    uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= *n_tasks) return;
    const uint32_t t = order[pos];
    uint32_t e = task_start[t];
    const uint32_t end = e + task_len[t];
    ZG_CHECK(t < *n_tasks && end <= p.n_entries);
    FqSlots S;
    vs_acc_inf(S);
    uint32_t absorbed = 0; // points added so far: the second one meets an affine accumulator (madd_aff: 6 products, not 10)
    auto fetch = [&](uint32_t &pc, FqSlots &S_) -> bool {
        if (e >= end) {
            vs_store_acc(partial + t, S_);
            return false;
        }
        for (;;) {
            uint32_t ent = entries[e++];
            ZG_CHECK((ent & 0x7fffffffu) < p.n_bases);
            const Affine *pt = bases + (ent & 0x7fffffffu);
            vs_set_stream(S_, VS_QX, &pt->x); // each base copy entry is read once per MSM: keep it out of the way of the slot files
            if (ent >> 31)
                vs_set_neg<true>(S_, VS_QY, &pt->y);
            else
                vs_set_stream(S_, VS_QY, &pt->y);
            // (0,0) encodes infinity in the affine image (only an expanded copy of a 2-torsion point)
            if (!(vs_is_zero(S_, VS_QX) && vs_is_zero(S_, VS_QY))) break;
            if (e >= end) {
                vs_store_acc(partial + t, S_);
                return false;
            }
        }
        pc = absorbed == 1 ? VPC_MADD_AFF : VPC_MADD_ACC;
        absorbed++;
        return true;
    };
    uint32_t pc = 0;
    if (!fetch(pc, S)) return;
#if ZG_ACC_SHARED_STASH
    fqvm_run_noinv(S, pc, [&](uint32_t, uint32_t &pc_, FqSlots &S_) { return fetch(pc_, S_); }, stash + threadIdx.x, 128);
#else
    fqvm_run_noinv(S, pc, [&](uint32_t, uint32_t &pc_, FqSlots &S_) { return fetch(pc_, S_); });
#endif
}

// ------------------------------------------------------------------ accumulate: batched affine pair tree
// Bucket b holds n_b entries at entries[off[b] .. off[b] + n_b). Round r halves every bucket:
// level r has cnt_r(b) = ceil(n_b / 2^r) affine points per bucket, level r+1 gets
//     out[i] = in[2i] + in[2i+1]   (or a copy of in[2i] when the count is odd),
// until one point per bucket is left. Level 0 is the base image addressed through `entries`
// (index | sign bit); level r >= 1 is a dense array of affine points where bucket b starts at
//     base_r(b) = (off[b] >> r) + b
// (closed form, buckets cannot overlap because floor((o+n)/2^r) - floor(o/2^r) + 1 >= ceil(n/2^r)),
// so level r has at most (M >> r) + n_buckets slots and levels alternate between two buffers.
// All additions of a round are independent: a thread owns K consecutive output slots and adds its
// K pairs in affine coordinates with ONE field inversion (Montgomery's trick over the
// denominators qx - px; programs tfwd/tinv/tbwd of the Fq VM): 6 products per addition +
// inversion / K, against 10 for the XYZZ mixed addition, and the work of a round divides evenly
// over threads whatever the bucket sizes (no tasks, no length sort, no heavy-bucket fold).
// The special cases of libff's add / mixed_add (bw6_761_g1.cpp:160-358: an operand at infinity,
// equal points -> doubling, opposite points -> infinity) are per-pair flags; infinity is (0,0).
// The point pair is read again from L2/HBM in the backward pass instead of being kept.
__device__ __forceinline__ uint32_t tree_rounds_needed(uint32_t max_count)
{
    uint32_t r = 0;
    while ((1u << r) < max_count) r++;
    return r;
}

// Rounds the tree runs for the largest bucket `max_count` when every bucket does at least `rmin` rounds
// and the chains that follow take at most MSM_TREE_CHAIN_MAX points: the level k_tree_chain reads.
__device__ __forceinline__ uint32_t tree_last_level(uint32_t max_count, uint32_t rmin)
{
    const uint32_t need = tree_rounds_needed(max_count);
    uint32_t r = rmin < need ? rmin : need;
    while (r < need && ((max_count + (1u << r) - 1) >> r) > MSM_TREE_CHAIN_MAX) r++;
    return r;
}

// level-r slot index of bucket b's first point (r >= 1)
__device__ __forceinline__ uint32_t tree_base(const uint32_t *off, uint32_t b, uint32_t r) { return (off[b] >> r) + b; }

template <class SF>
__device__ __forceinline__ void tree_load_entry(SF &S, uint32_t sx, uint32_t sy, const Affine *bases, uint32_t ent)
{
    const Affine *pt = bases + (ent & 0x7fffffffu);
    vs_set(S, sx, &pt->x);
    if (ent >> 31)
        vs_set_neg(S, sy, &pt->y);
    else
        vs_set(S, sy, &pt->y);
}

template <class SF>
__device__ __forceinline__ void tree_store_slots(Affine *dst, const SF &S, uint32_t sx, uint32_t sy)
{
    vs_get(&dst->x, S, sx);
    vs_get(&dst->y, S, sy);
}

__device__ __forceinline__ void affine_store_inf(Affine *dst)
{
    uint4 *d = reinterpret_cast<uint4 *>(dst);
#pragma unroll
    for (int k = 0; k < 12; k++) d[k] = make_uint4(0, 0, 0, 0);
}

template <int K>
__global__ void __launch_bounds__(128, 3)
k_tree_round(uint32_t r, uint32_t rmin, uint32_t n_buckets, const uint32_t *off, const TreeMeta *meta, const Affine *bases,
             const uint32_t *entries, const Affine *in, Affine *out)
{
    using SF = FqSlotFile<VS_COUNT + K>;
    if (r >= tree_last_level(meta->max_count, rmin)) return; // the chains take over from here
    const uint32_t slots = (off[n_buckets] >> (r + 1)) + n_buckets; // level r+1
    const uint32_t s0 = (blockIdx.x * blockDim.x + threadIdx.x) * K;
    // bucket of slot s0: the largest b with base_{r+1}(b) <= s0
    uint32_t b = 0, cur_base = 0, next_base = 0;
    if (s0 < slots) {
        uint32_t lo = 0, hi = n_buckets - 1;
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            if (tree_base(off, mid, r + 1) <= s0)
                lo = mid;
            else
                hi = mid - 1;
        }
        b = lo;
        cur_base = tree_base(off, b, r + 1);
        next_base = tree_base(off, b + 1, r + 1);
    }
    // Pass 1 (no arithmetic): classify the K slots. Padding slots are skipped, an odd point out and a
    // pair with an operand at infinity are copied; real additions are listed in pidx/oidx so that
    // the 32 lanes of the warp then run the same VM steps in lock-step whatever their slot mix.
    uint32_t pidx[K]; // level-r index of P (Q = next)
    uint8_t oidx[K];  // output slot - s0
    uint32_t na = 0;
    auto point_ptr = [&](uint32_t pos) -> const Affine * { return r == 0 ? bases + (entries[pos] & 0x7fffffffu) : in + pos; };
    auto quick_inf = [&](const Affine *pt) -> bool { // (0,0); the first 16 bytes decide in all but 2^-128 of cases
        const uint4 *w = reinterpret_cast<const uint4 *>(pt);
        uint4 v = w[0];
        if (v.x | v.y | v.z | v.w) return false;
        uint32_t acc = 0;
        for (int k = 1; k < 12; k++) {
            v = w[k];
            acc |= v.x | v.y | v.z | v.w;
        }
        return acc == 0;
    };
    auto copy_point = [&](Affine *dst, uint32_t pos) {
        const uint4 *w = reinterpret_cast<const uint4 *>(point_ptr(pos));
        uint4 *d = reinterpret_cast<uint4 *>(dst);
        const bool neg = r == 0 && (entries[pos] >> 31);
        if (!neg) {
#pragma unroll
            for (int k = 0; k < 12; k++) d[k] = w[k];
            return;
        }
#pragma unroll
        for (int k = 0; k < 6; k++) d[k] = w[k];
        uint32_t y[24], z[24], ny[24];
#pragma unroll
        for (int k = 0; k < 6; k++) {
            uint4 v = w[6 + k];
            y[4 * k] = v.x;
            y[4 * k + 1] = v.y;
            y[4 * k + 2] = v.z;
            y[4 * k + 3] = v.w;
            z[4 * k] = z[4 * k + 1] = z[4 * k + 2] = z[4 * k + 3] = 0;
        }
        fq_sub_regs(ny, z, y);
#pragma unroll
        for (int k = 0; k < 6; k++) d[6 + k] = make_uint4(ny[4 * k], ny[4 * k + 1], ny[4 * k + 2], ny[4 * k + 3]);
    };
    for (uint32_t j = 0; j < (uint32_t)K; j++) {
        const uint32_t s = s0 + j;
        if (s >= slots) break;
        while (s >= next_base) {
            b++;
            cur_base = next_base;
            next_base = tree_base(off, b + 1, r + 1);
        }
        const uint32_t i = s - cur_base;
        const uint32_t o = off[b], nb = off[b + 1] - o;
        const uint32_t cr = (nb + (1u << r) - 1) >> r; // points of this bucket at level r
        if (i >= ((cr + 1) >> 1)) continue;            // padding slot
        const uint32_t pos = (r == 0 ? o : (o >> r) + b) + 2 * i;
        if (2 * i + 1 >= cr) { // odd one out: carried up unchanged
            copy_point(out + s, pos);
            continue;
        }
        if (quick_inf(point_ptr(pos))) {
            copy_point(out + s, pos + 1);
            continue;
        }
        if (quick_inf(point_ptr(pos + 1))) {
            copy_point(out + s, pos);
            continue;
        }
        pidx[na] = pos;
        oidx[na] = (uint8_t)j;
        na++;
    }
    // Pass 2: forward products of the denominators (x coordinates only), one inversion, backward
    // additions. The lanes of a warp wait for each other before the inversion so that it is issued
    // once per warp. (Sharing one inversion among the four warps of the CTA through shared memory
    // was measured too: the three waiting warps cost more than the three inversions saved.)
    __syncwarp();
    if (na == 0) return;
    SF S;
    uint32_t act = 0, dbl = 0, rel = VS_COUNT, j = 0, phase = 0;
    auto load_x = [&](SF &S_, uint32_t jj) {
        rel = VS_COUNT + jj;
        vs_set(S_, VS_X, &point_ptr(pidx[jj])->x);
        vs_set(S_, VS_QX, &point_ptr(pidx[jj] + 1)->x);
    };
    auto load_y1 = [&](SF &S_, uint32_t sy, uint32_t pos) {
        const Affine *pt = point_ptr(pos);
        if (r == 0 && (entries[pos] >> 31))
            vs_set_neg(S_, sy, &pt->y);
        else
            vs_set(S_, sy, &pt->y);
    };
    auto load_pair = [&](SF &S_, uint32_t jj) {
        load_x(S_, jj);
        load_y1(S_, VS_Y, pidx[jj]);
        load_y1(S_, VS_QY, pidx[jj] + 1);
    };
    auto next_backward = [&](uint32_t &pc, SF &S_) -> bool {
        if (!act) return false;
        j = 31u - (uint32_t)__clz((int)act);
        load_pair(S_, j);
        pc = ((dbl >> j) & 1u) ? VPC_TBWD_DBL : VPC_TBWD;
        return true;
    };
    vs_one(S, VS_K1);
    load_x(S, 0);
    fqvm_run(
        S, VPC_TFWD,
        [&](uint32_t id, uint32_t &pc, SF &S_) -> bool {
            if (phase == 0) {
                if (id == 3) { // equal x: the special cases need the y coordinates
                    load_y1(S_, VS_Y, pidx[j]);
                    load_y1(S_, VS_QY, pidx[j] + 1);
                    return true;
                }
                if (id == 2) { // equal points: tangent, denominator 2 py already in T0
                    dbl |= 1u << j;
                    pc = VPC_TFWD_GO;
                    return true;
                }
                if (id == 0)
                    act |= 1u << j;
                else
                    affine_store_inf(out + s0 + oidx[j]); // P + (-P)
                if (++j < na) {
                    load_x(S_, j);
                    pc = VPC_TFWD;
                    return true;
                }
                __syncwarp(); // lanes with fewer additions wait here
                if (!act) return false;
                phase = 1;
                pc = VPC_TINV; // T4 = 1 / product
                return true;
            }
            if (phase == 2) {
                tree_store_slots(out + s0 + oidx[j], S_, VS_T2, VS_T3);
                act &= ~(1u << j);
            }
            phase = 2;
            return next_backward(pc, S_);
        },
        rel);
}

// What the tree rounds left of bucket b - cr = ceil(n_b / 2^R) affine points at level R = tree_last_level
// (the entries themselves when R = 0) - is summed by one XYZZ chain: partial[task_off[b]] = their sum.
__global__ void __launch_bounds__(128, 3)
k_tree_chain(uint32_t rmin, uint32_t n_buckets, const uint32_t *off, const uint32_t *task_off, const TreeMeta *meta,
             const Affine *bases, const uint32_t *entries, const Affine *lvl_odd, const Affine *lvl_even, XYZZ *partial)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_buckets) return;
    const uint32_t o = off[b], nb = off[b + 1] - o;
    if (nb == 0) return;
    const uint32_t R = tree_last_level(meta->max_count, rmin);
    const uint32_t cr = (nb + (1u << R) - 1) >> R;
    const Affine *lvl = R == 0 ? nullptr : ((R & 1u) ? lvl_odd : lvl_even) + tree_base(off, b, R);
    XYZZ *dst = partial + task_off[b];
    uint32_t i = 0;
    FqSlots S;
    vs_acc_inf(S);
    auto fetch = [&](uint32_t &pc, FqSlots &S_) -> bool {
        for (;;) {
            if (i >= cr) {
                vs_store_acc(dst, S_);
                return false;
            }
            if (R == 0) {
                tree_load_entry(S_, VS_QX, VS_QY, bases, entries[o + i]);
            } else {
                vs_set(S_, VS_QX, &lvl[i].x);
                vs_set(S_, VS_QY, &lvl[i].y);
            }
            i++;
            if (!(vs_is_zero(S_, VS_QX) && vs_is_zero(S_, VS_QY))) break; // (0,0) = infinity
        }
        pc = VPC_MADD;
        return true;
    };
    uint32_t pc = 0;
    if (!fetch(pc, S)) return;
    fqvm_run_noinv(S, pc, [&](uint32_t, uint32_t &pc_, FqSlots &S_) { return fetch(pc_, S_); });
}

// ------------------------------------------------------------------ tail: sum_b (b+1) B_b, batched over MSMs
// The second half of an MSM turns the bucket sums B_b (b < NB = 2^(c-1), per bucket set v) into
//     R_v = sum_b (b + 1) B_b = Root_v + sum_j 2^j S_{j,v},   S_j = sum of the buckets whose index has bit j set,
// and the result sum_v 2^(c v) R_v into libff special form. libff does this with a serial running sum per
// window (multiexp_accumulate_buckets, multiexp.tcc:90-125: 2 NB dependent additions) and c doublings
// between windows (:614-616). A dependent point addition costs ~50 us on one warp (14 products of ~3.5 us
// dependent-issue latency each), so the tail is latency-bound and is organised for DEPTH, one point
// operation per thread per launch, every launch serving all MSMs of a proof (TailBatch):
//   k_tail_fold   : buckets that were cut into several accumulate tasks are folded into their first partial
//                   (one warp per bucket, pair tree; nothing to do for ordinary buckets)
//   k_tail_level  : launch k = 1 .. c-1: X_k[s] = X_{k-1}[2s] + X_{k-1}[2s+1] (X_0 = buckets), and, riding
//                   along, the bit trees: Y_j starts as the odd items of X_j and is halved by every later
//                   launch, so that all S_j are complete after launch c-2 and the root after launch c-1.
//                   Work 2 NB additions (the running sum's), depth c-1.
//   k_tail_scale  : D_{v,j} = 2^(j + c v) S_{j,v} (j + c v doublings, in parallel), D_{v,c-1} = 2^(c v) Root_v
//   k_tail_sum    : pair-tree sum of all D (one CTA per MSM)
//   k_tail_norm   : to_special (one inversion per MSM, all MSMs of the batch side by side)
// Critical path for a pinned image (V = 1): c-1 additions + c-2 doublings + log2(c) additions + one
// inversion, about 2 ms, against ~75 dependent additions + the inversion (4.6 ms) for the radix-4 blocked
// running sum with a one-thread Horner that it replaces.
static constexpr uint32_t TAIL_MAX_JOBS = 8;
static constexpr uint32_t TAIL_FOLD_BLOCKS = 148 * 2; // warps per MSM folding cut buckets

struct TailJob {
    const uint32_t *task_off;    // [V*NB + 1]: bucket b's partial sums are partial[task_off[b] .. task_off[b+1])
    const uint32_t *heavy_list;  // buckets cut into > 1 task ...
    const uint32_t *heavy_count; // ... and how many
    XYZZ *partial;
    XYZZ *lvl[2];                // level-k buffer = lvl[k & 1]: [X_k | Y_0 | ... | Y_{k-1}]
    XYZZ *fin;                   // n_fin terms to sum (V * c scaled terms of one MSM, or the per-device partial sums of a sharded one)
    Proj *result;                // nullptr: leave the un-normalised sum in fin[0] (a shard of a multi-device MSM)
    uint32_t cm1;                // c - 1 = log2(NB)
    uint32_t V;
    uint32_t n_fin;
    uint32_t lvl_cap, partial_cap, fin_cap, n_buckets; // capacities in items (checked build)
};

struct TailBatch {
    uint32_t n_jobs;
    uint32_t first_block[TAIL_MAX_JOBS + 1]; // CTA range of each job in the current launch
    TailJob job[TAIL_MAX_JOBS];
};

__device__ __forceinline__ uint32_t tail_find_job(const TailBatch &tb, uint32_t blk)
{
    uint32_t j = 0;
    while (j + 1 < tb.n_jobs && blk >= tb.first_block[j + 1]) j++;
    return j;
}

// threads of launch k for one job
__host__ __device__ __forceinline__ uint32_t tail_level_threads(uint32_t cm1, uint32_t V, uint32_t k)
{
    if (k > cm1) return 0;
    const uint32_t nk = (1u << cm1) >> k;
    return V * nk + (k < cm1 ? k * V * (nk >> 1) : 0);
}

__global__ void __launch_bounds__(32, 1) k_tail_fold(TailBatch tb)
{
    const uint32_t j = tail_find_job(tb, blockIdx.x);
    const TailJob &J = tb.job[j];
    const uint32_t n_heavy = *J.heavy_count, nblk = tb.first_block[j + 1] - tb.first_block[j];
    for (uint32_t h = blockIdx.x - tb.first_block[j]; h < n_heavy; h += nblk) {
        const uint32_t b = J.heavy_list[h];
        ZG_CHECK(b < J.n_buckets);
        const uint32_t t0 = J.task_off[b], nt = J.task_off[b + 1] - t0;
        ZG_CHECK(nt > 1 && t0 + nt <= J.partial_cap);
        for (uint32_t step = 1; step < nt; step <<= 1) {
            // pairs (i, i + step), i = 2*step*q, q < ceil((nt - step) / (2 step))
            const uint32_t n_pairs = (nt - step + 2 * step - 1) / (2 * step);
            for (uint32_t q = threadIdx.x; q < n_pairs; q += 32) {
                XYZZ *dst = J.partial + t0 + 2 * step * q;
                FqSlots S;
                vs_load_acc(S, dst);
                vs_load_q(S, dst + step);
                fqvm_run_noinv(S, VPC_ADD, [&](uint32_t, uint32_t &, FqSlots &S_) {
                    vs_store_acc(dst, S_);
                    return false;
                });
            }
            __threadfence_block();
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(128, 3) k_tail_level(TailBatch tb, uint32_t k)
{
    const uint32_t j = tail_find_job(tb, blockIdx.x);
    const TailJob &J = tb.job[j];
    const uint32_t tid = (blockIdx.x - tb.first_block[j]) * blockDim.x + threadIdx.x;
    if (tid >= tail_level_threads(J.cm1, J.V, k)) return;
    const uint32_t nk = (1u << J.cm1) >> k, n_main = J.V * nk, n_tree = J.V * (nk >> 1);
    const XYZZ *in = J.lvl[(k - 1) & 1];
    XYZZ *out = J.lvl[k & 1];
    // operand items a, b (nullptr = infinity) and the output slot
    const XYZZ *a, *b;
    XYZZ *dst;
    auto x_prev = [&](uint32_t i) -> const XYZZ * { // item i of X_{k-1}
        if (k > 1) {
            ZG_CHECK(i < J.lvl_cap);
            return in + i;
        }
        ZG_CHECK(i < J.n_buckets);
        const uint32_t t0 = J.task_off[i];
        return t0 == J.task_off[i + 1] ? nullptr : J.partial + t0;
    };
    if (tid < n_main) {
        a = x_prev(2 * tid);
        b = x_prev(2 * tid + 1);
        dst = out + tid;
    } else {
        const uint32_t t = tid - n_main, tj = t / n_tree, u = t - tj * n_tree;
        if (tj == k - 1) { // first halving of tree k-1: the odd items of X_{k-1}
            a = x_prev(4 * u + 1);
            b = x_prev(4 * u + 3);
        } else {           // Y_tj of the previous launch: 2 n_tree items behind the 2 n_main of X_{k-1}
            const XYZZ *y = in + 2 * n_main + (size_t)tj * (2 * n_tree);
            ZG_CHECK(2 * n_main + (size_t)tj * (2 * n_tree) + 2 * u + 1 < J.lvl_cap);
            a = y + 2 * u;
            b = y + 2 * u + 1;
        }
        dst = out + n_main + (size_t)tj * n_tree + u;
    }
    ZG_CHECK((size_t)(dst - out) < J.lvl_cap);
    FqSlots S;
    if (a)
        vs_load_acc(S, a);
    else
        vs_acc_inf(S);
    if (b)
        vs_load_q(S, b);
    else
        vs_zero(S, VS_QZZ); // the add program only tests QZZ
    fqvm_run_noinv(S, VPC_ADD, [&](uint32_t, uint32_t &, FqSlots &S_) {
        vs_store_acc(dst, S_);
        return false;
    });
}

// fin[v*c + j] = 2^(j + c v) S_{j,v} for j < c-1, fin[v*c + c-1] = 2^(c v) Root_v
__global__ void __launch_bounds__(128, 3) k_tail_scale(TailBatch tb)
{
    const uint32_t jb = tail_find_job(tb, blockIdx.x);
    const TailJob &J = tb.job[jb];
    const uint32_t tid = (blockIdx.x - tb.first_block[jb]) * blockDim.x + threadIdx.x;
    const uint32_t c = J.cm1 + 1, V = J.V;
    if (tid >= V * c) return;
    const uint32_t v = tid / c, j = tid - v * c;
    const XYZZ *s2 = J.lvl[(J.cm1 - 1) & 1]; // level c-2: X (2 per set) | Y_0 .. Y_{c-3} (1 per set)
    const XYZZ *src;
    uint32_t ndbl = c * v;
    if (j == J.cm1) {
        src = J.lvl[J.cm1 & 1] + v; // root
    } else {
        src = j + 1 == J.cm1 ? s2 + 2 * v + 1 : s2 + 2 * V + (size_t)j * V + v;
        ndbl += j;
    }
    ZG_CHECK((size_t)(src - J.lvl[0]) < J.lvl_cap || (size_t)(src - J.lvl[1]) < J.lvl_cap);
    ZG_CHECK(tid < J.fin_cap);
    FqSlots S;
    vs_load_acc(S, src);
    XYZZ *dst = J.fin + tid;
    if (ndbl == 0 || vs_is_zero(S, VS_ZZ)) {
        vs_store_acc(dst, S);
        return;
    }
    fqvm_run_noinv(S, VPC_DBL, [&](uint32_t, uint32_t &pc, FqSlots &S_) {
        if (--ndbl) {
            pc = VPC_DBL;
            return true;
        }
        vs_store_acc(dst, S_);
        return false;
    });
}

// fin[0] = sum of the job's V*c terms: pair tree, one CTA per job
__global__ void __launch_bounds__(128, 1) k_tail_sum(TailBatch tb)
{
    const TailJob &J = tb.job[blockIdx.x];
    const uint32_t n = J.n_fin;
    for (uint32_t step = 1; step < n; step <<= 1) {
        const uint32_t n_pairs = (n - step + 2 * step - 1) / (2 * step);
        for (uint32_t q = threadIdx.x; q < n_pairs; q += blockDim.x) {
            XYZZ *dst = J.fin + 2 * step * q;
            FqSlots S;
            vs_load_acc(S, dst);
            vs_load_q(S, dst + step);
            fqvm_run_noinv(S, VPC_ADD, [&](uint32_t, uint32_t &, FqSlots &S_) {
                vs_store_acc(dst, S_);
                return false;
            });
        }
        __threadfence_block();
        __syncthreads();
    }
}

// result = fin[0] in libff special form (to_special, bw6_761_g1.cpp:58-77); thread i serves job i
__global__ void __launch_bounds__(32, 1) k_tail_norm(TailBatch tb)
{
    if (threadIdx.x >= tb.n_jobs) return;
    const TailJob &J = tb.job[threadIdx.x];
    Proj *out = J.result;
    if (!out) return;
    FqSlots S;
    vs_load_acc(S, J.fin);
    fqvm_run(S, VPC_TOAFF, [&](uint32_t id, uint32_t &, FqSlots &S_) {
        vs_one(S_, VS_T0);
        if (id == 2) { // infinity: (0, one, 0)
            uint4 *d = reinterpret_cast<uint4 *>(out);
            for (int q = 0; q < 18; q++) d[q] = make_uint4(0, 0, 0, 0);
            vs_get(&out->Y, S_, VS_T0);
        } else {
            vs_get(&out->X, S_, VS_QX);
            vs_get(&out->Y, S_, VS_QY);
            vs_get(&out->Z, S_, VS_T0);
        }
        return false;
    });
}

// ------------------------------------------------------------------ fixed-base batch exponentiation
// out[i] = (coeff * s_i) * g for one base g and n scalars: libff::batch_exp / batch_exp_with_coeff
// (multiexp.tcc:808-947) followed by batch_to_special (:949-974), the work of the trusted setup
// (r1cs_gg_ppzksnark_generator, r1cs_gg_ppzksnark.tcc:330-390; kc_batch_exp, kc_multiexp.tcc:115-239).
// The reference builds a table of all window values on one thread and walks it per scalar with
// general additions; here:
//   k_bexp_powers : B_w = 2^(c w) g, one thread per window (c w doublings)          -> XYZZ
//   k_bexp_table  : T[w][j-1] = j B_w, j < 2^c, one thread per entry (double-and-add) -> XYZZ
//   k_xyzz_to_affine after each (one inversion per 8 points, Montgomery's trick)
//   k_bexp_eval   : one thread per scalar: sum_w T[w][digit_w(s)] as a chain of mixed additions
//   k_xyzz_to_affine -> libff special form {x, y, one} / zero {0, one, 0}
// Window choice is free (the result is a group element, compared in normalised form).
static constexpr uint32_t BEXP_C = 12;
static constexpr uint32_t BEXP_W = (FR_SCALAR_BITS + BEXP_C - 1) / BEXP_C; // 32 windows
static constexpr uint32_t BEXP_ROW = (1u << BEXP_C) - 1;                   // table entries per window
static constexpr uint32_t TOAFF_K = 8;

// in[i] XYZZ (zz == 0: infinity) -> affine (0,0 = infinity) and/or libff special form.
__global__ void __launch_bounds__(128, 2)
k_xyzz_to_affine(const XYZZ *in, uint32_t n, Affine *out_aff, Proj *out_proj, Fq *scratch_prefix)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i0 = t * TOAFF_K;
    if (i0 >= n) return;
    const uint32_t i1 = i0 + TOAFF_K < n ? i0 + TOAFF_K : n;
    FqSlots S;
    uint32_t i = i0, stage = 0; // 0 forward products, 1 inverting, 2 backward
    vs_one(S, VS_K1);
    auto write_inf = [&](uint32_t k) {
        if (out_aff) affine_store_inf(out_aff + k);
        if (out_proj) {
            uint4 *d = reinterpret_cast<uint4 *>(out_proj + k);
            for (int q = 0; q < 18; q++) d[q] = make_uint4(0, 0, 0, 0);
            for (int q = 0; q < 6; q++) d[6 + q] = make_uint4(FQ_ONE[4 * q], FQ_ONE[4 * q + 1], FQ_ONE[4 * q + 2], FQ_ONE[4 * q + 3]);
        }
    };
    auto next_forward = [&](uint32_t &pc, FqSlots &S_) -> bool {
        while (i < i1) {
            vs_load_acc(S_, in + i);
            if (!vs_is_zero(S_, VS_ZZ)) {
                vs_get(scratch_prefix + i, S_, VS_K1);
                pc = VPC_PRODSTEP; // K1 *= ZZZ
                return true;
            }
            i++;
        }
        return false;
    };
    auto next_backward = [&](uint32_t &pc, FqSlots &S_) -> bool {
        while (i > i0) {
            i--;
            vs_load_acc(S_, in + i);
            if (!vs_is_zero(S_, VS_ZZ)) {
                vs_set(S_, VS_T3, scratch_prefix + i);
                pc = VPC_UNSTEP;
                return true;
            }
            write_inf(i);
        }
        return false;
    };
    uint32_t pc0 = 0;
    if (!next_forward(pc0, S)) { // all at infinity
        for (uint32_t k = i0; k < i1; k++) write_inf(k);
        return;
    }
    fqvm_run(S, pc0, [&](uint32_t, uint32_t &pc, FqSlots &S_) {
        if (stage == 0) {
            i++;
            if (next_forward(pc, S_)) return true;
            stage = 1;
            pc = VPC_INVPROD; // T0 = 1 / K1
            return true;
        }
        if (stage == 1) {
            stage = 2;
            i = i1;
            return next_backward(pc, S_);
        }
        if (out_aff) {
            vs_get(&out_aff[i].x, S_, VS_QX);
            vs_get(&out_aff[i].y, S_, VS_QY);
        }
        if (out_proj) {
            Proj *o = out_proj + i;
            vs_get(&o->X, S_, VS_QX);
            vs_get(&o->Y, S_, VS_QY);
            uint4 *d = reinterpret_cast<uint4 *>(&o->Z);
            for (int q = 0; q < 6; q++) d[q] = make_uint4(FQ_ONE[4 * q], FQ_ONE[4 * q + 1], FQ_ONE[4 * q + 2], FQ_ONE[4 * q + 3]);
        }
        return next_backward(pc, S_);
    });
}

// B_w = 2^(c w) g for w < W: thread w doubles g (affine, finite) c*w times.
__global__ void __launch_bounds__(32, 1) k_bexp_powers(const Affine *g, XYZZ *out)
{
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= BEXP_W) return;
    FqSlots S;
    vs_set(S, VS_X, &g->x);
    vs_set(S, VS_Y, &g->y);
    vs_one(S, VS_ZZ);
    vs_one(S, VS_ZZZ);
    uint32_t left = BEXP_C * w;
    if (left == 0) {
        vs_store_acc(out + w, S);
        return;
    }
    fqvm_run_noinv(S, VPC_DBL, [&](uint32_t, uint32_t &pc, FqSlots &S_) {
        if (--left) {
            pc = VPC_DBL;
            return true;
        }
        vs_store_acc(out + w, S_);
        return false;
    });
}

// T[w][j-1] = j * B_w (B_w affine; (0,0) if the base has small order), 1 <= j <= BEXP_ROW:
// double-and-add over the bits of j below its top bit.
__global__ void __launch_bounds__(128, 2) k_bexp_table(const Affine *powers, XYZZ *out)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= BEXP_W * BEXP_ROW) return;
    const uint32_t w = t / BEXP_ROW, j = t - w * BEXP_ROW + 1;
    FqSlots S;
    const Affine *b = powers + w;
    vs_set(S, VS_QX, &b->x);
    vs_set(S, VS_QY, &b->y);
    if (vs_is_zero(S, VS_QX) && vs_is_zero(S, VS_QY)) {
        xyzz_store_inf(out + t);
        return;
    }
    vs_set(S, VS_X, &b->x);
    vs_set(S, VS_Y, &b->y);
    vs_one(S, VS_ZZ);
    vs_one(S, VS_ZZZ);
    int bit = 30 - __clz((int)j); // next bit below the top one
    if (bit < 0) {
        vs_store_acc(out + t, S);
        return;
    }
    uint32_t phase = 0; // 0: a doubling just ran, 1: an addition just ran
    fqvm_run_noinv(S, VPC_DBL, [&](uint32_t, uint32_t &pc, FqSlots &S_) {
        if (phase == 0 && ((j >> bit) & 1u)) {
            phase = 1;
            pc = VPC_MADD;
            return true;
        }
        phase = 0;
        if (bit == 0) {
            vs_store_acc(out + t, S_);
            return false;
        }
        bit--;
        pc = VPC_DBL;
        return true;
    });
}

// out[i] = sum_w T[w][digit_w - 1], digits = BEXP_C-bit windows of the plain integer coeff * s_i.
__global__ void __launch_bounds__(128, 3)
k_bexp_eval(const Affine *table, const uint8_t *scalars, size_t sstride, uint32_t n, const Fr *coeff, XYZZ *out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    __align__(8) uint32_t v[14];
    {
        uint32_t m[12], t[12];
        fr_load(m, reinterpret_cast<const Fr *>(scalars + (size_t)i * sstride));
        if (coeff) {
            uint32_t cf[12];
            fr_load(cf, coeff);
            fr_mul_regs(t, m, cf);
#pragma unroll
            for (int k = 0; k < 12; k++) m[k] = t[k];
        }
        fr_from_mont(t, m);
#pragma unroll
        for (int k = 0; k < 12; k++) v[k] = t[k];
        v[12] = v[13] = 0;
    }
    FqSlots S;
    vs_acc_inf(S);
    uint32_t w = 0;
    auto fetch = [&](uint32_t &pc, FqSlots &S_) -> bool {
        for (; w < BEXP_W; w++) {
            const uint32_t bit = w * BEXP_C, lo = bit >> 5, sh = bit & 31u;
            const uint64_t two = (uint64_t)v[lo] | ((uint64_t)v[lo + 1] << 32);
            const uint32_t d = (uint32_t)(two >> sh) & BEXP_ROW;
            if (!d) continue;
            const Affine *pt = table + (size_t)w * BEXP_ROW + (d - 1);
            vs_set(S_, VS_QX, &pt->x);
            vs_set(S_, VS_QY, &pt->y);
            if (vs_is_zero(S_, VS_QX) && vs_is_zero(S_, VS_QY)) continue; // table entry at infinity
            w++;
            pc = VPC_MADD;
            return true;
        }
        vs_store_acc(out + i, S_);
        return false;
    };
    uint32_t pc0 = 0;
    if (!fetch(pc0, S)) return;
    fqvm_run_noinv(S, pc0, [&](uint32_t, uint32_t &pc, FqSlots &S_) { return fetch(pc, S_); });
}

// ------------------------------------------------------------------ workspace
struct MsmBuffers {
    int32_t *digits;
    uint32_t *counts, *offsets, *cursor, *ntasks, *task_off, *task_start, *task_len, *task_bucket, *order, *len_hist, *len_cursor,
        *entries, *scan_tmp;
    XYZZ *partial, *lvl[2], *fin;
    Affine *tree_lvl[2]; // pair-tree levels: odd levels in [0], even levels >= 2 in [1]
    TreeMeta *tree_meta;
    Proj *result;
    size_t total;
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

inline MsmBuffers msm_layout(const MsmPlan &p, uint8_t *base)
{
    MsmBuffers b;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        uint8_t *ptr = base ? base + off : nullptr;
        off += align_up(bytes, 256);
        return ptr;
    };
    size_t nd = (size_t)p.n * p.W;
    size_t nbk = p.n_buckets;
    b.digits = (int32_t *)take(nd * 4);
    b.counts = (uint32_t *)take((nbk + 1) * 4);
    b.offsets = (uint32_t *)take((nbk + 1) * 4);
    b.cursor = (uint32_t *)take((nbk + 1) * 4);
    b.ntasks = (uint32_t *)take((nbk + 1) * 4);
    b.task_off = (uint32_t *)take((nbk + 1) * 4);
    b.task_start = (uint32_t *)take(p.max_tasks * 4);
    b.task_len = (uint32_t *)take(p.max_tasks * 4);
    b.task_bucket = (uint32_t *)take(p.max_tasks * 4);
    b.order = (uint32_t *)take(p.max_tasks * 4);
    b.len_hist = (uint32_t *)take((MSM_LEN_BINS + 2) * 4);
    b.len_cursor = (uint32_t *)take((MSM_LEN_BINS + 2) * 4);
    b.entries = (uint32_t *)take(nd * 4);
    b.scan_tmp = (uint32_t *)take((nbk / (SCAN_BLOCK * SCAN_ITEMS) + 2) * 4);
    b.partial = (XYZZ *)take(p.max_tasks * sizeof(XYZZ));
    // tail level buffers: launch k writes V * (n_k + k * n_k / 2) items, largest at k = 1 (3/4 of the buckets)
    for (int i = 0; i < 2; i++) b.lvl[i] = (XYZZ *)take((nbk - nbk / 4) * sizeof(XYZZ));
    b.fin = (XYZZ *)take((size_t)p.V * p.c * sizeof(XYZZ));
    b.tree_lvl[0] = b.tree_lvl[1] = nullptr;
    b.tree_meta = (TreeMeta *)take(sizeof(TreeMeta));
    if (p.acc_mode == MSM_ACC_TREE) {
        b.tree_lvl[0] = (Affine *)take(((nd >> 1) + nbk + 1) * sizeof(Affine));
        b.tree_lvl[1] = (Affine *)take(((nd >> 2) + nbk + 1) * sizeof(Affine));
    }
    b.result = (Proj *)take(sizeof(Proj));
    b.total = off;
    return b;
}

// Number of kernel launches msm_launch_front issues for plan p (the tail counts its own).
inline uint64_t msm_front_launch_count(const MsmPlan &p)
{
    if (p.acc_mode == MSM_ACC_TREE) return 1 + 3 + 1 + 1 + 3 + p.tree_rounds + 1;
    return 1 + 3 + 1 + 1 + 3 + 1 + 3 + 1 + 1;
}

// K pairs per thread for a round expected to hold `pairs` additions: a larger K spreads the
// inversion thinner, a smaller one keeps a full wave of threads (148 SM x 3 CTA x 128) in the air.
template <int K>
inline void tree_round_launch(const MsmPlan &p, const MsmBuffers &b, uint32_t r, const Affine *d_bases, cudaStream_t st)
{
    const uint64_t slots = (((uint64_t)p.n * p.W) >> (r + 1)) + p.n_buckets;
    const unsigned blocks = (unsigned)(((slots + K - 1) / K + 127) / 128);
    k_tree_round<K><<<blocks, 128, 0, st>>>(r, p.tree_min, (uint32_t)p.n_buckets, b.offsets, b.tree_meta, d_bases, b.entries,
                                            r ? b.tree_lvl[(r - 1) & 1] : nullptr, b.tree_lvl[r & 1]);
}

// First half of an MSM on `st`: digit decomposition, counting sort by bucket, task cutting and
// the bucket accumulation kernel (the throughput-bound part). d_bases: affine image (G copies of
// copy_stride points); d_scalars: Fr Montgomery.
inline cudaError_t msm_launch_front(const MsmPlan &p, const MsmBuffers &b, const Affine *d_bases, const uint8_t *d_is_inf,
                                    const uint8_t *d_scalars, size_t sstride, const uint64_t *d_index, cudaStream_t st,
                                    cudaEvent_t acc_ev0 = nullptr, cudaEvent_t acc_ev1 = nullptr)
{
    const uint32_t n = p.n;
    cudaMemsetAsync(b.counts, 0, (p.n_buckets + 1) * 4, st);
    cudaMemsetAsync(b.len_hist, 0, (MSM_LEN_BINS + 2) * 4, st);
    k_digits<<<(n + 127) / 128, 128, 0, st>>>(d_scalars, sstride, d_index, d_is_inf, p, b.digits, b.counts);
    exclusive_scan_u32(b.counts, p.n_buckets, b.offsets, b.scan_tmp, st);
    cudaMemcpyAsync(b.cursor, b.offsets, (p.n_buckets + 1) * 4, cudaMemcpyDeviceToDevice, st);
    size_t nd = (size_t)n * p.W;
    k_scatter<<<(unsigned)((nd + 255) / 256), 256, 0, st>>>(p, b.digits, b.cursor, b.entries);
    unsigned gb = (unsigned)((p.n_buckets + 255) / 256);
    if (p.acc_mode == MSM_ACC_TREE) {
        cudaMemsetAsync(b.tree_meta, 0, sizeof(TreeMeta), st);
        k_tree_count<<<gb, 256, 0, st>>>(p, b.counts, b.ntasks, b.tree_meta);
        exclusive_scan_u32(b.ntasks, p.n_buckets, b.task_off, b.scan_tmp, st);
        if (acc_ev0) cudaEventRecord(acc_ev0, st);
        const uint64_t wave = 148ull * 3 * 128;
        for (uint32_t r = 0; r < p.tree_rounds; r++) {
            const uint64_t pairs = ((uint64_t)n * p.W) >> (r + 1);
            static const char *force_k = getenv("ZECALE_GPU_TREE_K"); // experiments: 8 | 16 | 32 for every round
            const uint32_t fk = force_k ? (uint32_t)atoi(force_k) : 0; // 0 = by size
            if (fk ? fk == 32 : pairs >= 32 * wave)
                tree_round_launch<32>(p, b, r, d_bases, st);
            else if (fk ? fk == 16 : pairs >= 16 * wave)
                tree_round_launch<16>(p, b, r, d_bases, st);
            else
                tree_round_launch<8>(p, b, r, d_bases, st);
        }
        k_tree_chain<<<(unsigned)((p.n_buckets + 127) / 128), 128, 0, st>>>(p.tree_min, (uint32_t)p.n_buckets, b.offsets, b.task_off,
                                                                           b.tree_meta, d_bases, b.entries, b.tree_lvl[0],
                                                                           b.tree_lvl[1], b.partial);
        if (acc_ev1) cudaEventRecord(acc_ev1, st);
        return cudaGetLastError();
    }
    k_task_count<<<gb, 256, 0, st>>>(p, b.counts, b.ntasks);
    exclusive_scan_u32(b.ntasks, p.n_buckets, b.task_off, b.scan_tmp, st);
    // heavy-bucket list: reuses ntasks (dead after its scan); counter = spare word of len_hist
    k_task_fill<<<gb, 256, 0, st>>>(p, b.counts, b.offsets, b.task_off, b.task_start, b.task_len, b.task_bucket, b.len_hist,
                                    b.ntasks, b.len_hist + MSM_LEN_BINS + 1);
    exclusive_scan_u32(b.len_hist, MSM_LEN_BINS + 1, b.len_cursor, b.scan_tmp, st);
    const uint32_t *n_tasks = b.task_off + p.n_buckets;
    k_task_order<<<(unsigned)((p.max_tasks + 255) / 256), 256, 0, st>>>(n_tasks, b.task_len, b.len_cursor, b.order);
    if (acc_ev0) cudaEventRecord(acc_ev0, st);
    k_accumulate<<<(unsigned)((p.max_tasks + 127) / 128), 128, 0, st>>>(p, d_bases, n_tasks, b.order, b.task_start,
                                                                         b.task_len, b.entries, b.partial);
    if (acc_ev1) cudaEventRecord(acc_ev1, st);
    return cudaGetLastError();
}

// Second half of up to TAIL_MAX_JOBS MSMs whose bucket accumulation is done (or ordered before this
// on `st`): see the tail section above. Each result (libff special form, 288 B) is left in its
// buffers.result. Returns the number of launches through *n_launches.
inline TailJob msm_tail_job(const MsmPlan &p, const MsmBuffers &b)
{
    TailJob j{};
    j.task_off = b.task_off;
    j.heavy_list = b.ntasks; // reused by k_task_fill once the task scan has consumed it
    j.heavy_count = b.len_hist + MSM_LEN_BINS + 1;
    j.partial = b.partial;
    j.lvl[0] = b.lvl[0];
    j.lvl[1] = b.lvl[1];
    j.fin = b.fin;
    j.result = b.result;
    j.cm1 = p.c - 1;
    j.V = p.V;
    j.n_fin = p.V * p.c;
    j.n_buckets = (uint32_t)p.n_buckets;
    j.lvl_cap = (uint32_t)(p.n_buckets - p.n_buckets / 4);
    j.partial_cap = (uint32_t)p.max_tasks;
    j.fin_cap = p.V * p.c;
    return j;
}

// Last two steps only (sum n_fin terms, normalise): combines the per-device partial sums of sharded MSMs.
inline cudaError_t msm_launch_gather_batch(const TailJob *jobs, uint32_t n_jobs, cudaStream_t st)
{
    if (n_jobs == 0 || n_jobs > TAIL_MAX_JOBS) return cudaErrorInvalidValue;
    TailBatch tb{};
    tb.n_jobs = n_jobs;
    for (uint32_t i = 0; i < n_jobs; i++) tb.job[i] = jobs[i];
    k_tail_sum<<<n_jobs, 128, 0, st>>>(tb);
    k_tail_norm<<<1, 32, 0, st>>>(tb);
    return cudaGetLastError();
}

inline cudaError_t msm_launch_tail_batch(const TailJob *jobs, uint32_t n_jobs, cudaStream_t st, uint64_t *n_launches = nullptr)
{
    if (n_jobs == 0 || n_jobs > TAIL_MAX_JOBS) return cudaErrorInvalidValue;
    // the level ping-pong below needs at least two levels (level 0 lives in `partial`, k_tail_scale reads the last two
    // of lvl[]): window sizes outside [MSM_MIN_C, MSM_MAX_C] are a caller error, not a plan
    for (uint32_t i = 0; i < n_jobs; i++)
        if (jobs[i].cm1 + 1 < MSM_MIN_C || jobs[i].cm1 + 1 > MSM_MAX_C) return cudaErrorInvalidValue;
    TailBatch tb{};
    tb.n_jobs = n_jobs;
    uint32_t max_cm1 = 0;
    for (uint32_t i = 0; i < n_jobs; i++) {
        tb.job[i] = jobs[i];
        if (jobs[i].cm1 > max_cm1) max_cm1 = jobs[i].cm1;
    }
    uint64_t nl = 0;
    // blocks(i) CTAs of `threads` per job
    auto ranges = [&](auto blocks) -> uint32_t {
        uint32_t acc = 0;
        for (uint32_t i = 0; i < n_jobs; i++) {
            tb.first_block[i] = acc;
            acc += blocks(tb.job[i]);
        }
        for (uint32_t i = n_jobs; i <= TAIL_MAX_JOBS; i++) tb.first_block[i] = acc;
        return acc;
    };
    uint32_t g = ranges([](const TailJob &) { return TAIL_FOLD_BLOCKS; });
    k_tail_fold<<<g, 32, 0, st>>>(tb);
    nl++;
    for (uint32_t k = 1; k <= max_cm1; k++) {
        g = ranges([k](const TailJob &J) { return (tail_level_threads(J.cm1, J.V, k) + 127) / 128; });
        k_tail_level<<<g, 128, 0, st>>>(tb, k);
        nl++;
    }
    g = ranges([](const TailJob &J) { return (J.V * (J.cm1 + 1) + 127) / 128; });
    k_tail_scale<<<g, 128, 0, st>>>(tb);
    k_tail_sum<<<n_jobs, 128, 0, st>>>(tb);
    k_tail_norm<<<1, 32, 0, st>>>(tb);
    nl += 3;
    if (n_launches) *n_launches = nl;
    return cudaGetLastError();
}

} // namespace zecale
