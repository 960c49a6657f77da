// Pippenger bucket MSM over BW6-761 G1/G2 for sm_100a.
//
// Replaces libff::multi_exp<G, Fr, multi_exp_method_BDLO12_signed, special>
// (libff/algebra/scalar_multiplication/multiexp.tcc:507-688) and its callers
// multi_exp_filter_one_zero (:695-757) / kc_multi_exp_with_mixed_addition
// (libsnark knowledge_commitment/kc_multiexp.tcc:31-113).
//
// The reference processes one c-bit window at a time, top-down, on one OpenMP
// chunk per thread. Here all windows are processed at once:
//   1. k_digits      scalar -> plain integer (as_bigint) -> signed c-bit digits
//                    (field_get_signed_digit semantics, field_utils.tcc:81-115),
//                    histogram of (window, |digit|) with global atomics
//   2. scan          exclusive prefix sum -> bucket offsets
//   3. k_scatter     counting sort of (point, sign) into bucket order
//   4. k_tasks_*     buckets larger than S entries are split into even sub-tasks
//   5. k_accumulate  one thread per task: serial mixed additions into an XYZZ sum
//   6. k_reduce1     per (window, segment of L buckets): fold task partials and run
//                    the running-sum trick (multiexp_accumulate_buckets, :90-125)
//   7. k_reduce2     per window: combine segments (suffix scan + tree), no serial pass
//   8. k_final       Horner over windows (c doublings each, multiexp.tcc:612-629),
//                    normalise to libff special form
// No 0/1 filter is needed: a zero scalar has all-zero digits, a one is an ordinary
// bucket entry, and base points at infinity are skipped by the mixed addition.
#pragma once
#include <cstdint>
#include <cstdio>

#include "ec.cuh"

namespace zecale {

struct MsmPlan {
    uint32_t n;        // points
    uint32_t c;        // window bits
    uint32_t W;        // windows
    uint32_t NB;       // buckets per window = 2^(c-1)
    uint32_t S;        // max entries per accumulate task
    uint32_t L;        // buckets per reduce1 segment
    uint32_t nseg;     // NB / L
    uint64_t n_buckets;// W * NB
    uint64_t max_tasks;
};

static constexpr int FR_SCALAR_BITS = 377;

inline MsmPlan msm_make_plan(uint32_t n, int force_c = 0)
{
    MsmPlan p;
    p.n = n;
    uint32_t lg = 0;
    while ((1ull << (lg + 1)) <= n) lg++;
    int c = (int)lg - 3;
    if (c < 4) c = 4;
    if (c > 16) c = 16;
    if (force_c) c = force_c;
    p.c = c;
    p.W = (FR_SCALAR_BITS + 1 + c - 1) / c;
    p.NB = 1u << (c - 1);
    uint32_t avg = (n + p.NB - 1) / p.NB;
    p.S = avg * 2 < 16 ? 16 : avg * 2;
    p.L = p.NB >= 16 ? 16 : p.NB;
    p.nseg = p.NB / p.L;
    p.n_buckets = (uint64_t)p.W * p.NB;
    p.max_tasks = ((uint64_t)n * p.W) / p.S + p.n_buckets + 1;
    return p;
}

// ------------------------------------------------------------------ import
// libff special-form points (stride bytes apart, {X,Y,Z}) -> device affine image.
// status bit0 is raised if a base is neither Z == one nor Z == 0.
__global__ void k_import_special(const uint8_t *raw, size_t stride, size_t offset, uint32_t n, Affine *out, int *status)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 *src = reinterpret_cast<const uint4 *>(raw + (size_t)i * stride + offset);
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
        return;
    }
    if (zne != 0) atomicOr(status, 1);
#pragma unroll
    for (int k = 0; k < 12; k++) dst[k] = src[k];
}

// ------------------------------------------------------------------ digits
// One thread per scalar. scalars are Montgomery Fr, `sstride` bytes apart; the scalar of
// base i is scalars[index ? index[i] : i].
__global__ void k_digits(const uint8_t *scalars, size_t sstride, const uint64_t *index, MsmPlan p, int32_t *digits, uint32_t *counts)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    size_t si = index ? (size_t)index[i] : (size_t)i;
    uint32_t m[12], v[13];
    fr_load(m, reinterpret_cast<const Fr *>(scalars + si * sstride));
    {
        uint32_t t[12];
        fr_from_mont(t, m);
#pragma unroll
        for (int k = 0; k < 12; k++) v[k] = t[k];
        v[12] = 0;
    }
    const uint32_t c = p.c;
    const uint32_t mask = (1u << c) - 1;
    const uint32_t half = 1u << (c - 1);
    uint32_t carry = 0;
    for (uint32_t w = 0; w < p.W; w++) {
        uint32_t bit = w * c;
        uint32_t lo = bit >> 5, sh = bit & 31;
        // v is indexed dynamically: keep it in local memory, 13 words
        uint64_t two = (uint64_t)v[lo] | ((uint64_t)(lo + 1 < 13 ? v[lo + 1] : 0u) << 32);
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
        digits[(size_t)w * p.n + i] = sd;
        if (sd != 0) {
            uint32_t b = (sd < 0 ? (uint32_t)(-sd) : (uint32_t)sd) - 1;
            atomicAdd(&counts[(size_t)w * p.NB + b], 1u);
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
    uint32_t pos = atomicAdd(&cursor[(size_t)w * p.NB + b], 1u);
    entries[pos] = i | (neg << 31);
}

// ------------------------------------------------------------------ scan (uint32, exclusive)
// 3-phase: per-block reduce, single-block scan of block sums, per-block scan + offset.
static constexpr int SCAN_BLOCK = 1024;
static constexpr int SCAN_ITEMS = 4; // per thread

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *smem, uint32_t &total)
{
    // Hillis-Steele in shared memory over blockDim.x (<= 1024) values
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

// in-place exclusive scan of block_sums (nb <= SCAN_BLOCK*SCAN_ITEMS*... handled by loop)
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
__global__ void k_task_count(MsmPlan p, const uint32_t *counts, uint32_t *ntasks)
{
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.n_buckets) return;
    uint32_t cnt = counts[b];
    ntasks[b] = (cnt + p.S - 1) / p.S;
}

__global__ void k_task_fill(MsmPlan p, const uint32_t *task_off, uint32_t *task_bucket)
{
    size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.n_buckets) return;
    uint32_t t0 = task_off[b], t1 = task_off[b + 1];
    for (uint32_t t = t0; t < t1; t++) task_bucket[t] = (uint32_t)b;
}

// ------------------------------------------------------------------ accumulate
__global__ void __launch_bounds__(128, 4)
k_accumulate(MsmPlan p, const Affine *bases, const uint32_t *counts, const uint32_t *offsets,
             const uint32_t *task_off, const uint32_t *task_bucket, const uint32_t *entries, XYZZ *partial)
{
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t T = task_off[p.n_buckets];
    if (t >= T) return;
    uint32_t b = task_bucket[t];
    uint32_t t0 = task_off[b];
    uint32_t nt = task_off[b + 1] - t0;
    uint32_t k = t - t0;
    uint64_t cnt = counts[b];
    uint32_t lo = offsets[b] + (uint32_t)((cnt * k) / nt);
    uint32_t hi = offsets[b] + (uint32_t)((cnt * (k + 1)) / nt);
    XYZZ acc;
    xyzz_set_inf(acc);
    for (uint32_t e = lo; e < hi; e++) {
        uint32_t ent = entries[e];
        xyzz_madd(acc, bases[ent & 0x7fffffffu], (ent >> 31) != 0);
    }
    xyzz_copy(partial[t], acc);
}

// ------------------------------------------------------------------ reduce
// One thread per (window, segment of L buckets): bucket value = sum of its task
// partials; running sum gives seg_sum = sum B_i and seg_acc = sum (i+1) B_i.
__global__ void __launch_bounds__(128, 4)
k_reduce1(MsmPlan p, const uint32_t *task_off, const XYZZ *partial, XYZZ *seg_sum, XYZZ *seg_acc)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; // global segment id = w*nseg + s
    if (g >= p.W * p.nseg) return;
    size_t b0 = (size_t)g * p.L;
    XYZZ run, acc;
    xyzz_set_inf(run);
    xyzz_set_inf(acc);
    for (int i = (int)p.L - 1; i >= 0; i--) {
        uint32_t t0 = task_off[b0 + i], t1 = task_off[b0 + i + 1];
        for (uint32_t t = t0; t < t1; t++) xyzz_add(run, partial[t]);
        xyzz_add(acc, run);
    }
    xyzz_copy(seg_sum[g], run);
    xyzz_copy(seg_acc[g], acc);
}

// One block per window. window_sum = sum_s seg_acc[s] + L * sum_s s*seg_sum[s];
// sum_s s*S_s = sum_{j>=1} suffix_j  (suffix scan, Hillis-Steele, ping-pong in global).
__global__ void __launch_bounds__(256, 1)
k_reduce2(MsmPlan p, XYZZ *seg_sum, XYZZ *seg_acc, XYZZ *scratch, XYZZ *window_sum)
{
    const uint32_t w = blockIdx.x;
    const uint32_t ns = p.nseg;
    XYZZ *A = seg_acc + (size_t)w * ns;
    XYZZ *Y0 = seg_sum + (size_t)w * ns;
    XYZZ *Y1 = scratch + (size_t)w * ns;
    XYZZ tmp;
    // suffix scan of seg_sum
    for (uint32_t d = 1; d < ns; d <<= 1) {
        for (uint32_t i = threadIdx.x; i < ns; i += blockDim.x) {
            xyzz_copy(tmp, Y0[i]);
            if (i + d < ns) xyzz_add(tmp, Y0[i + d]);
            xyzz_copy(Y1[i], tmp);
        }
        __syncthreads();
        XYZZ *t = Y0;
        Y0 = Y1;
        Y1 = t;
    }
    // Y0[i] = sum_{j>=i} S_j. Drop i = 0 (weight 0) by overwriting it with infinity.
    if (threadIdx.x == 0) xyzz_set_inf(Y0[0]);
    __syncthreads();
    // tree sums of Y0[0..ns) and A[0..ns)
    for (uint32_t d = 1; d < ns; d <<= 1) {
        for (uint32_t i = threadIdx.x * 2 * d; i + d < ns; i += blockDim.x * 2 * d) {
            xyzz_copy(tmp, Y0[i]);
            xyzz_add(tmp, Y0[i + d]);
            xyzz_copy(Y0[i], tmp);
            xyzz_copy(tmp, A[i]);
            xyzz_add(tmp, A[i + d]);
            xyzz_copy(A[i], tmp);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        xyzz_copy(tmp, Y0[0]);
        for (uint32_t l = 1; l < p.L; l <<= 1) xyzz_dbl(tmp);
        xyzz_add(tmp, A[0]);
        xyzz_copy(window_sum[w], tmp);
    }
}

// Horner over the windows, top-down, then normalise. One thread.
__global__ void k_final(MsmPlan p, const XYZZ *window_sum, Proj *out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    XYZZ r;
    xyzz_copy(r, window_sum[p.W - 1]);
    for (int w = (int)p.W - 2; w >= 0; w--) {
        for (uint32_t k = 0; k < p.c; k++) xyzz_dbl(r);
        xyzz_add(r, window_sum[w]);
    }
    xyzz_to_special(*out, r);
}

// ------------------------------------------------------------------ workspace
struct MsmWorkspace {
    void *buf = nullptr;
    size_t cap = 0;
};

struct MsmBuffers {
    int32_t *digits;
    uint32_t *counts, *offsets, *cursor, *ntasks, *task_off, *task_bucket, *entries, *scan_tmp;
    XYZZ *partial, *seg_sum, *seg_acc, *scratch, *window_sum;
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
    b.task_bucket = (uint32_t *)take(p.max_tasks * 4);
    b.entries = (uint32_t *)take(nd * 4);
    b.scan_tmp = (uint32_t *)take((nbk / (SCAN_BLOCK * SCAN_ITEMS) + 2) * 4);
    b.partial = (XYZZ *)take(p.max_tasks * sizeof(XYZZ));
    size_t nsegs = (size_t)p.W * p.nseg;
    b.seg_sum = (XYZZ *)take(nsegs * sizeof(XYZZ));
    b.seg_acc = (XYZZ *)take(nsegs * sizeof(XYZZ));
    b.scratch = (XYZZ *)take(nsegs * sizeof(XYZZ));
    b.window_sum = (XYZZ *)take((size_t)p.W * sizeof(XYZZ));
    b.result = (Proj *)take(sizeof(Proj));
    b.total = off;
    return b;
}

// Launch the whole MSM on `st`. d_bases: affine image; d_scalars: Fr Montgomery.
// Result (libff special form, 288 B) is left in buffers.result on the device.
inline cudaError_t msm_launch(const MsmPlan &p, const MsmBuffers &b, const Affine *d_bases, const uint8_t *d_scalars,
                              size_t sstride, const uint64_t *d_index, cudaStream_t st)
{
    const uint32_t n = p.n;
    cudaMemsetAsync(b.counts, 0, (p.n_buckets + 1) * 4, st);
    k_digits<<<(n + 127) / 128, 128, 0, st>>>(d_scalars, sstride, d_index, p, b.digits, b.counts);
    exclusive_scan_u32(b.counts, p.n_buckets, b.offsets, b.scan_tmp, st);
    cudaMemcpyAsync(b.cursor, b.offsets, (p.n_buckets + 1) * 4, cudaMemcpyDeviceToDevice, st);
    size_t nd = (size_t)n * p.W;
    k_scatter<<<(unsigned)((nd + 255) / 256), 256, 0, st>>>(p, b.digits, b.cursor, b.entries);
    unsigned gb = (unsigned)((p.n_buckets + 255) / 256);
    k_task_count<<<gb, 256, 0, st>>>(p, b.counts, b.ntasks);
    exclusive_scan_u32(b.ntasks, p.n_buckets, b.task_off, b.scan_tmp, st);
    k_task_fill<<<gb, 256, 0, st>>>(p, b.task_off, b.task_bucket);
    k_accumulate<<<(unsigned)((p.max_tasks + 127) / 128), 128, 0, st>>>(p, d_bases, b.counts, b.offsets, b.task_off,
                                                                         b.task_bucket, b.entries, b.partial);
    uint32_t nsegs = p.W * p.nseg;
    k_reduce1<<<(nsegs + 127) / 128, 128, 0, st>>>(p, b.task_off, b.partial, b.seg_sum, b.seg_acc);
    k_reduce2<<<p.W, 256, 0, st>>>(p, b.seg_sum, b.seg_acc, b.scratch, b.window_sum);
    k_final<<<1, 32, 0, st>>>(p, b.window_sum, b.result);
    return cudaGetLastError();
}

} // namespace zecale
