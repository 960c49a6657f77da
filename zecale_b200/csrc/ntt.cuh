// Radix-2 NTT over BW6-761 Fr (377 bit) for sm_100a.
//
// Replaces libfqfft::basic_radix2_domain<bw6_761_Fr>::{FFT, iFFT, cosetFFT, icosetFFT}
// (libfqfft/evaluation_domain/domains/basic_radix2_domain.tcc:55-88) and the kernels
// behind them (_basic_serial_radix2_FFT / _basic_parallel_radix2_FFT,
// basic_radix2_domain_aux.tcc:44-172; _multiply_by_coset :174-183). Natural order in
// and out, Montgomery form in and out; field arithmetic is exact so any correct
// factorisation gives the reference's bits.
//
// Algorithm: multi-step ("4-step") Cooley-Tukey. m = R1*R2(*R3); each pass runs
// batches of 2^logR-point DIT sub-transforms entirely in shared memory (bit-reversed
// on load, natural on store), multiplies by the inter-step twiddle w_m^(u*k) on the
// way out and writes through an affine index map, so no separate transpose or
// bit-reversal pass ever touches HBM: 2 passes (<= 2^22) or 3 passes = 192*m / 288*m
// bytes of traffic. The CPU analogue is the reference's 2-step OpenMP split
// (basic_radix2_domain_aux.tcc:105-147).
//
// Twiddles: one table w_m^k, k < m/2, per domain size, built once per context
// (48*m/2 bytes; at 2^24 that is 403 MB of the 180 GB HBM). Inverse transforms use
// w^-k = -w^(m/2-k) from the same table.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <utility>
#include <vector>

#include "fqvm.cuh"

namespace zecale {

// ------------------------------------------------------------------ Fr helpers
__device__ __forceinline__ void fr_neg_regs(uint32_t (&r)[12], const uint32_t (&a)[12])
{
    uint32_t z[12];
#pragma unroll
    for (int k = 0; k < 12; k++) z[k] = 0;
    fr_sub_regs(r, z, a);
}

// r = a^e for a 64-bit exponent (square and multiply, MSB first).
__device__ __forceinline__ void fr_pow_u64(uint32_t (&r)[12], const uint32_t (&a)[12], uint64_t e)
{
    uint32_t acc[12];
    fr_set_const(acc, FR_ONE);
    for (int bit = 63; bit >= 0; bit--) {
        fr_sqr_regs(acc, acc);
        if ((e >> bit) & 1) fr_mul_regs(acc, acc, a);
    }
#pragma unroll
    for (int k = 0; k < 12; k++) r[k] = acc[k];
}

// r = a^(r-2) (Fermat inverse); 0 -> 0.
__device__ __forceinline__ void fr_inv(uint32_t (&r)[12], const uint32_t (&a)[12])
{
    uint32_t acc[12];
    fr_set_const(acc, FR_ONE);
    for (int bit = FR_BITS - 1; bit >= 0; bit--) {
        fr_sqr_regs(acc, acc);
        if ((FR_PM2[bit >> 5] >> (bit & 31)) & 1u) fr_mul_regs(acc, acc, a);
    }
#pragma unroll
    for (int k = 0; k < 12; k++) r[k] = acc[k];
}

// table[i] = scale * base^i, i < n. Thread t seeds base^(t*chunk) by fast
// exponentiation and walks `chunk` consecutive powers.
__global__ void k_fr_powers(Fr base, Fr scale, Fr *table, size_t n, uint32_t chunk)
{
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t i0 = t * chunk;
    if (i0 >= n) return;
    uint32_t b[12], cur[12], s[12];
    fr_load(b, &base);
    fr_load(s, &scale);
    fr_pow_u64(cur, b, (uint64_t)i0);
    fr_mul_regs(cur, cur, s);
    for (uint32_t k = 0; k < chunk && i0 + k < n; k++) {
        fr_store(table + i0 + k, cur);
        fr_mul_regs(cur, cur, b);
    }
}

// a[i] *= table[i]
__global__ void k_fr_scale_table(Fr *a, const Fr *table, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t x[12], y[12];
    fr_load(x, a + i);
    fr_load(y, table + i);
    fr_mul_regs(x, x, y);
    fr_store(a + i, x);
}

// h[i] = (a[i]*b[i] - c[i]) * zinv   — the pointwise core of r1cs_to_qap_witness_map
// (libsnark r1cs_to_qap.tcc:283-315: H_tmp = aA∘aB, -= aC, divide_by_Z_on_coset). In place in a.
__global__ void k_qap_pointwise(Fr *a, const Fr *b, const Fr *c, const Fr *zinv, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t x[12], y[12], z[12];
    fr_load(x, a + i);
    fr_load(y, b + i);
    fr_mul_regs(x, x, y);
    fr_load(y, c + i);
    fr_sub_regs(x, x, y);
    fr_load(z, zinv);
    fr_mul_regs(x, x, z);
    fr_store(a + i, x);
}

// Evaluation vectors of r1cs_to_qap_witness_map (libsnark r1cs_to_qap.tcc:236-248, 293-296) from a CSR
// image of one constraint matrix: out[i] = <M_i, w> for i < n_rows (linear_combination::evaluate,
// relations/variable.tcc:288: sum of coeff * assignment[index], index 0 = the constant one, which
// is w[0] here), and for the A matrix the input-consistency rows out[n_rows + i] = w[i], i <= n_inputs
// (:236-240); everything else up to the domain size m is zero. One thread per row (rows hold 1-3 terms).
__global__ void k_r1cs_eval(const uint64_t *row_ptr, const uint32_t *col, const Fr *val, const Fr *w, uint32_t n_rows,
                            uint32_t n_inputs_plus_one /* 0 for B and C */, Fr *out, size_t m)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    uint32_t acc[12];
#pragma unroll
    for (int k = 0; k < 12; k++) acc[k] = 0;
    if (i < n_rows) {
        for (uint64_t t = row_ptr[i]; t < row_ptr[i + 1]; t++) {
            uint32_t c[12], x[12];
            fr_load(c, val + t);
            fr_load(x, w + col[t]);
            fr_mul_regs(c, c, x);
            fr_add_regs(acc, acc, c);
        }
    } else if (i - n_rows < n_inputs_plus_one) {
        fr_load(acc, w + (i - n_rows));
    }
    fr_store(out + i, acc);
}

// out = (g^(2^log_m) - 1)^-1 : the vanishing polynomial on the coset g*H is the constant
// g^m - 1 (libfqfft basic_radix2_domain.tcc:103-126 compute_vanishing_polynomial / divide_by_Z_on_coset).
__global__ void k_fr_zinv(Fr *out, Fr g, uint32_t log_m)
{
    uint32_t x[12], one[12];
    fr_load(x, &g);
    for (uint32_t i = 0; i < log_m; i++) fr_sqr_regs(x, x);
    fr_set_const(one, FR_ONE);
    fr_sub_regs(x, x, one);
    fr_inv(x, x);
    fr_store(out, x);
}

// ------------------------------------------------------------------ pass kernel
struct NttPass {
    uint32_t logR;      // sub-transform length R = 2^logR
    uint32_t T;         // vectors per tile (power of two)
    uint32_t n_batch;   // vectors are indexed (beta < n_batch, u < n_inner)
    uint32_t n_inner;
    uint32_t tile_over_batch; // 1: a tile holds T consecutive beta (same u); 0: T consecutive u
    uint64_t ibs, isv, isr;   // in  index = beta*ibs + u*isv + r*isr
    uint64_t obs, osv, osr;   // out index = beta*obs + u*osv + k*osr
    uint32_t in_fast_v;       // 1: adjacent vectors are adjacent in memory on load (else adjacent r)
    uint32_t tw_post;         // 1: multiply output (u, k) by w_m^(u*k*tw_mul)
    uint64_t tw_mul;
    uint32_t log_m;           // full domain size (twiddle table has m/2 entries)
    uint32_t inverse;         // use w^-1
    uint64_t sub_tw_stride;   // w_R^e = w_m^(e*sub_tw_stride)
    // distributed (4-step across ranks) extensions; all zero for a single-GPU transform
    uint32_t r_lo_bits;       // != 0: input element r is split, address += (r >> r_lo_bits)*isr_hi + (r & lo)*isr
    uint64_t isr_hi;
    uint32_t k_lo_bits;       // != 0: output element k is split the same way with osr_hi
    uint64_t osr_hi;
    uint32_t tw_use_beta;     // inter-step twiddle exponent uses beta (row vectors) instead of u
    uint64_t tw_vec_off;      // global offset of this rank's vectors in the twiddle exponent
    const Fr *scale_tab;      // != null: output (natural index n) is also multiplied by scale_tab[n], where
    uint64_t sc_bs, sc_sv, sc_sr; // n = beta*sc_bs + u*sc_sv + k*sc_sr + sc_off (global index of the output)
    uint64_t sc_off;
    const Fr *in_scale_tab;   // != null: input element is first multiplied by in_scale_tab[u*isc_sv + r*isc_sr + isc_off]
    uint64_t isc_sv, isc_sr, isc_off;
    uint32_t stw_global;      // 1: sub-transform twiddles are read from the global table (R too large to stage)
    uint64_t batch_stride;    // blockIdx.y selects one of several same-sized vectors, batch_stride elements apart
};

__device__ __forceinline__ uint32_t bitrev(uint32_t x, uint32_t bits) { return __brev(x) >> (32 - bits); }

// w_m^e (or w_m^-e) from the half table tw[k] = w_m^k, k < m/2. e < m.
__device__ __forceinline__ void ntt_twiddle(uint32_t (&w)[12], const Fr *tw, uint64_t e, uint32_t log_m, uint32_t inverse)
{
    const uint64_t half = 1ull << (log_m - 1);
    bool neg = false;
    if (inverse && e != 0) { // w^-e = w^(m-e)
        e = (half << 1) - e;
    }
    if (e >= half) {
        e -= half;
        neg = true;
    }
    fr_load(w, tw + e);
    if (neg) fr_neg_regs(w, w);
}

extern __shared__ __align__(16) uint8_t ntt_smem[];

// Shared-memory layout of a tile. An Fr is 48 B = three 16-byte chunks, LDS.128 / STS.128 serve a quarter
// warp per wavefront, so 8 consecutive lanes are conflict-free iff their element positions are distinct
// mod 8 (3 is odd). Element i of a vector sits at position ntt_swz(i): the low three bits are XORed with
// GF(2)-linear images of bit 3 (-> 111) and of the top three bits (-> 001, 010, 100), which makes every
// access pattern of the kernel a bijection on the low three bits: butterflies of stage 0 / 1 / 2 (lanes vary
// bits {1,2,3} / {0,2,3} / {0,1,3}; columns 001, 010, 100, 111 are independent in threes), later stages (bits
// {0,1,2}), and the bit-reversed load (lanes vary the top bits). Vectors are padded so that the T vectors a
// quarter warp touches in the load / store phases fall on different residues as well.
__host__ __device__ __forceinline__ uint32_t ntt_vstride(uint32_t logR, uint32_t T)
{
    const uint32_t R = 1u << logR;
    return R + (T >= 8 ? 1u : T > 1 ? 8u / T : 0u);
}
// Sub-transform twiddle e (w_R^e, e < R/2) sits at e with its low three bits XORed by every higher 3-bit group of e:
// stage s reads e = j * 2^t for 8 consecutive j, i.e. varies bits t, t+1, t+2 of e, which the fold sends to three
// different low bits whatever t is - 8 different residues mod 8 instead of 8 entries a power of two apart.
__device__ __forceinline__ uint32_t ntt_swz_tw(uint32_t e)
{
    uint32_t f = e >> 3;
    f ^= f >> 3;
    f ^= f >> 6;   // R/2 <= 2^11: four groups
    return e ^ (f & 7u);
}
__device__ __forceinline__ uint32_t ntt_swz(uint32_t i, uint32_t logR)
{
    uint32_t x = i ^ (((i >> 3) & 1u) * 7u);
    if (logR >= 7) x ^= __brev(i >> (logR - 3)) >> 29;
    return x;
}

// resident CTAs per SM of k_ntt_pass (register cap 65536 / (256 * ZG_NTT_CTAS))
#ifndef ZG_NTT_CTAS
#define ZG_NTT_CTAS 3
#endif
__global__ void __launch_bounds__(256, ZG_NTT_CTAS) k_ntt_pass(const Fr *in, Fr *out, const Fr *__restrict__ tw, NttPass p)
{
    const uint32_t R = 1u << p.logR;
    const uint32_t T = p.T;
    const uint32_t vstride = ntt_vstride(p.logR, T);
    Fr *data = reinterpret_cast<Fr *>(ntt_smem);
    Fr *stw = data + (size_t)T * vstride; // sub-transform twiddles w_R^e, e < R/2, at position ntt_swz_tw(e)
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    in += (size_t)blockIdx.y * p.batch_stride;
    out += (size_t)blockIdx.y * p.batch_stride;

    // tile -> (beta0, u0)
    uint32_t beta0, u0;
    if (p.tile_over_batch) {
        uint32_t tiles_per_u = p.n_batch / T;
        u0 = blockIdx.x / tiles_per_u;
        beta0 = (blockIdx.x % tiles_per_u) * T;
    } else {
        uint32_t tiles_per_b = p.n_inner / T;
        beta0 = blockIdx.x / tiles_per_b;
        u0 = (blockIdx.x % tiles_per_b) * T;
    }

    // sub-transform twiddles w_R^e, e < R/2
    for (uint32_t e = tid; e < R / 2 && !p.stw_global; e += nthr) {
        uint32_t w[12];
        ntt_twiddle(w, tw, (uint64_t)e * p.sub_tw_stride, p.log_m, p.inverse);
        fr_store(stw + ntt_swz_tw(e), w);
    }
    // load, bit-reversing the position inside each vector
    for (uint32_t l = tid; l < T * R; l += nthr) {
        uint32_t v, r;
        if (p.in_fast_v) {
            v = l & (T - 1);
            r = l / T;
        } else {
            r = l & (R - 1);
            v = l >> p.logR;
        }
        uint32_t beta = beta0 + (p.tile_over_batch ? v : 0);
        uint32_t u = u0 + (p.tile_over_batch ? 0 : v);
        uint64_t idx = (uint64_t)beta * p.ibs + (uint64_t)u * p.isv;
        if (p.r_lo_bits)
            idx += (uint64_t)(r >> p.r_lo_bits) * p.isr_hi + (uint64_t)(r & ((1u << p.r_lo_bits) - 1)) * p.isr;
        else
            idx += (uint64_t)r * p.isr;
        uint32_t x[12];
        fr_load(x, in + idx);
        if (p.in_scale_tab) {
            uint32_t w[12];
            fr_load(w, p.in_scale_tab + ((uint64_t)u * p.isc_sv + (uint64_t)r * p.isc_sr + p.isc_off));
            fr_mul_regs(x, x, w);
        }
        fr_store(data + (size_t)v * vstride + ntt_swz(bitrev(r, p.logR), p.logR), x);
    }
    __syncthreads();
    // DIT stages
    const uint32_t halfR = R >> 1;
    for (uint32_t s = 0; s < p.logR; s++) {
        const uint32_t half = 1u << s;
        const uint32_t tw_step = halfR >> s; // w_(2half)^j = w_R^(j * R/(2half))
        for (uint32_t l = tid; l < T * halfR; l += nthr) {
            uint32_t v = l >> (p.logR - 1);
            uint32_t b = l & (halfR - 1);
            uint32_t j = b & (half - 1);
            uint32_t i = ((b >> s) << (s + 1)) + j;
            Fr *pa = data + (size_t)v * vstride + ntt_swz(i, p.logR);
            Fr *pb = data + (size_t)v * vstride + ntt_swz(i + half, p.logR);
            uint32_t x[12], y[12], w[12];
            fr_load(y, pb);
            if (j != 0) {
                if (p.stw_global)
                    ntt_twiddle(w, tw, (uint64_t)j * tw_step * p.sub_tw_stride, p.log_m, p.inverse);
                else
                    fr_load(w, stw + ntt_swz_tw(j * tw_step));
                fr_mul_regs(y, y, w);
            }
            fr_load(x, pa);
            fr_add_regs(w, x, y);
            fr_sub_regs(y, x, y);
            fr_store(pa, w);
            fr_store(pb, y);
        }
        __syncthreads();
    }
    // store (+ inter-step twiddle)
    for (uint32_t l = tid; l < T * R; l += nthr) {
        uint32_t v = l & (T - 1);
        uint32_t k = l / T;
        uint32_t beta = beta0 + (p.tile_over_batch ? v : 0);
        uint32_t u = u0 + (p.tile_over_batch ? 0 : v);
        uint32_t x[12];
        fr_load(x, data + (size_t)v * vstride + ntt_swz(k, p.logR));
        const uint64_t tv = (p.tw_use_beta ? beta : u) + p.tw_vec_off;
        if (p.tw_post && tv != 0 && k != 0) {
            uint32_t w[12];
            uint64_t e = (tv * k * p.tw_mul) & ((1ull << p.log_m) - 1);
            ntt_twiddle(w, tw, e, p.log_m, p.inverse);
            fr_mul_regs(x, x, w);
        }
        if (p.scale_tab) {
            uint32_t w[12];
            fr_load(w, p.scale_tab + ((uint64_t)beta * p.sc_bs + (uint64_t)u * p.sc_sv + (uint64_t)k * p.sc_sr + p.sc_off));
            fr_mul_regs(x, x, w);
        }
        uint64_t idx = (uint64_t)beta * p.obs + (uint64_t)u * p.osv;
        if (p.k_lo_bits)
            idx += (uint64_t)(k >> p.k_lo_bits) * p.osr_hi + (uint64_t)(k & ((1u << p.k_lo_bits) - 1)) * p.osr;
        else
            idx += (uint64_t)k * p.osr;
        fr_store(out + idx, x);
    }
}

// ------------------------------------------------------------------ tables
struct NttTables {
    std::map<uint32_t, Fr *> tw;                                   // log_m -> w_m^k, k < m/2
    std::map<std::pair<uint32_t, std::vector<uint32_t>>, Fr *> coset; // (log_m | inv<<8, g limbs) -> g^(+-i) (* 1/m)
    Fr *minv = nullptr;                                            // scratch for constants
};

inline void ntt_tables_free(NttTables &t)
{
    for (auto &kv : t.tw) cudaFree(kv.second);
    for (auto &kv : t.coset) cudaFree(kv.second);
    t.tw.clear();
    t.coset.clear();
}

// single-thread helper kernels for table seeds
__global__ void k_fr_root(Fr *out, uint32_t log_m, uint32_t inverse)
{
    // w_m = root_of_unity^(2^(s - log_m)) (libff get_root_of_unity, field_utils.tcc:190-206)
    uint32_t w[12];
    fr_set_const(w, inverse ? FR_ROOT_2S_INV : FR_ROOT_2S);
    for (uint32_t i = log_m; i < (uint32_t)FR_TWO_ADICITY; i++) fr_sqr_regs(w, w);
    fr_store(out, w);
}

// out[0] = g or g^-1 ; out[1] = 1 or (2^log_m)^-1  (Montgomery form)
__global__ void k_fr_coset_seed(Fr *out, Fr g, uint32_t log_m, uint32_t inverse)
{
    uint32_t x[12], s[12];
    fr_load(x, &g);
    fr_set_const(s, FR_ONE);
    if (inverse) {
        fr_inv(x, x);
        uint32_t two[12];
        fr_add_regs(two, s, s);
        uint32_t mm[12];
        fr_set_const(mm, FR_ONE);
        for (uint32_t i = 0; i < log_m; i++) fr_mul_regs(mm, mm, two);
        fr_inv(s, mm);
    }
    fr_store(out, x);
    fr_store(out + 1, s);
}

inline cudaError_t ntt_get_twiddles(NttTables &t, uint32_t log_m, cudaStream_t st, const Fr **out, unsigned *launches)
{
    auto it = t.tw.find(log_m);
    if (it != t.tw.end()) {
        *out = it->second;
        return cudaSuccess;
    }
    size_t half = (size_t)1 << (log_m - 1);
    Fr *d = nullptr, *seed = nullptr;
    cudaError_t e = cudaMalloc((void **)&d, half * sizeof(Fr));
    if (e != cudaSuccess) return e;
    e = cudaMalloc((void **)&seed, 2 * sizeof(Fr));
    if (e != cudaSuccess) return e;
    k_fr_root<<<1, 1, 0, st>>>(seed, log_m, 0);
    Fr hroot, hone;
    cudaMemcpyAsync(&hroot, seed, sizeof(Fr), cudaMemcpyDeviceToHost, st);
    cudaMemcpyFromSymbolAsync(&hone, FR_ONE, sizeof(Fr), 0, cudaMemcpyDeviceToHost, st);
    cudaStreamSynchronize(st);
    uint32_t chunk = 64;
    size_t nthreads = (half + chunk - 1) / chunk;
    k_fr_powers<<<(unsigned)((nthreads + 127) / 128), 128, 0, st>>>(hroot, hone, d, half, chunk);
    *launches += 2;
    e = cudaStreamSynchronize(st);
    cudaFree(seed);
    if (e != cudaSuccess) return e;
    t.tw[log_m] = d;
    *out = d;
    return cudaSuccess;
}

// table[i] = g^i (forward) or g^-i / m (inverse)
inline cudaError_t ntt_get_coset(NttTables &t, uint32_t log_m, const Fr &g, uint32_t inverse, cudaStream_t st,
                                 const Fr **out, unsigned *launches)
{
    std::vector<uint32_t> key(g.l, g.l + 12);
    auto k = std::make_pair(log_m | (inverse << 8), key);
    auto it = t.coset.find(k);
    if (it != t.coset.end()) {
        *out = it->second;
        return cudaSuccess;
    }
    size_t m = (size_t)1 << log_m;
    Fr *d = nullptr, *seed = nullptr;
    cudaError_t e = cudaMalloc((void **)&d, m * sizeof(Fr));
    if (e != cudaSuccess) return e;
    e = cudaMalloc((void **)&seed, 2 * sizeof(Fr));
    if (e != cudaSuccess) return e;
    k_fr_coset_seed<<<1, 1, 0, st>>>(seed, g, log_m, inverse);
    Fr h[2];
    cudaMemcpyAsync(h, seed, 2 * sizeof(Fr), cudaMemcpyDeviceToHost, st);
    cudaStreamSynchronize(st);
    uint32_t chunk = 64;
    size_t nthreads = (m + chunk - 1) / chunk;
    k_fr_powers<<<(unsigned)((nthreads + 127) / 128), 128, 0, st>>>(h[0], h[1], d, m, chunk);
    *launches += 2;
    e = cudaStreamSynchronize(st);
    cudaFree(seed);
    if (e != cudaSuccess) return e;
    t.coset[k] = d;
    *out = d;
    return cudaSuccess;
}

inline size_t ntt_pass_smem(uint32_t logR, uint32_t T)
{
    size_t R = (size_t)1 << logR;
    if (logR >= 12) return (size_t)T * ntt_vstride(logR, T) * sizeof(Fr); // twiddles stay in global memory (stw_global)
    return ((size_t)T * ntt_vstride(logR, T) + R / 2) * sizeof(Fr);
}

// Tile width: as many vectors as fit ~100 KB (2 CTAs / SM), at most 8.
inline uint32_t ntt_pick_T(uint32_t logR, uint32_t n_avail)
{
    uint32_t T = 8;
    while (T > 1 && ntt_pass_smem(logR, T) > 100 * 1024) T >>= 1;
    while (T > n_avail) T >>= 1;
    return T ? T : 1;
}

inline cudaError_t ntt_run_pass(const Fr *in, Fr *out, const Fr *tw, NttPass p, cudaStream_t st, uint32_t batch = 1)
{
    size_t smem = ntt_pass_smem(p.logR, p.T);
    p.stw_global = p.logR >= 12;
    cudaError_t e = cudaFuncSetAttribute(k_ntt_pass, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    uint64_t tiles = (uint64_t)p.n_batch * p.n_inner / p.T;
    k_ntt_pass<<<dim3((unsigned)tiles, batch), 256, smem, st>>>(in, out, tw, p);
    return cudaGetLastError();
}

// In-place (result back in `a`) transform of 2^log_m elements; `tmp` is a same-sized
// scratch vector. flags: bit0 inverse, bit1 coset (host pointer coset_g, Montgomery).
// batch > 1: the same transform of `batch` vectors stored back to back in a (and tmp), one launch per pass
// (the three independent A / B / C chains of r1cs_to_qap_witness_map).
inline cudaError_t ntt_launch(NttTables &tabs, Fr *a, Fr *tmp, uint32_t log_m, uint32_t flags, const Fr *coset_g,
                              cudaStream_t st, unsigned *launches, uint32_t batch = 1)
{
    const uint32_t inverse = flags & 1u, coset = (flags >> 1) & 1u;
    const size_t m = (size_t)1 << log_m;
    const Fr *tw = nullptr;
    cudaError_t e = ntt_get_twiddles(tabs, log_m, st, &tw, launches);
    if (e != cudaSuccess) return e;
    const Fr *ctab = nullptr;
    Fr one_g; // plain iFFT still needs the 1/m scaling: table for g = 1
    memcpy(one_g.l, FR_ONE_HOST, sizeof(Fr));
    if (coset || inverse) {
        e = ntt_get_coset(tabs, log_m, coset ? *coset_g : one_g, inverse, st, &ctab, launches);
        if (e != cudaSuccess) return e;
    }
    // cosetFFT: a[i] *= g^i before the transform (basic_radix2_domain.tcc:77-81) — fused into the
    // load of the first pass; iFFT: * 1/m, icosetFFT: additionally * g^-i (:63-88) — fused into
    // the store of the last pass. `first` / `last` set the table on the pass that needs it.
    const Fr *in_tab = (coset && !inverse) ? ctab : nullptr;
    const Fr *out_tab = inverse ? ctab : nullptr;
    NttPass p{};
    p.log_m = log_m;
    p.inverse = inverse;
    p.batch_stride = m;
    if (log_m <= 11) {
        p.logR = log_m;
        p.T = 1;
        p.n_batch = 1;
        p.n_inner = 1;
        p.tile_over_batch = 0;
        p.isr = 1;
        p.osr = 1;
        p.in_fast_v = 0;
        p.tw_post = 0;
        p.sub_tw_stride = 1;
        p.in_scale_tab = in_tab; p.isc_sr = 1;
        p.scale_tab = out_tab; p.sc_sr = 1;
        e = ntt_run_pass(a, tmp, tw, p, st, batch);
        if (e != cudaSuccess) return e;
        (*launches)++;
        e = cudaMemcpyAsync(a, tmp, (size_t)batch * m * sizeof(Fr), cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return e;
    } else if (log_m <= 22) {
        const uint32_t l1 = (log_m + 1) / 2, l2 = log_m - l1;
        const uint64_t m1 = 1ull << l1, m2 = 1ull << l2;
        // pass 1: columns. vector u = j2 < m2, element r = j1: idx = r*m2 + u
        p.logR = l1;
        p.n_batch = 1;
        p.n_inner = (uint32_t)m2;
        p.T = ntt_pick_T(l1, p.n_inner);
        p.tile_over_batch = 0;
        p.ibs = 0; p.isv = 1; p.isr = m2;
        p.obs = 0; p.osv = 1; p.osr = m2;
        p.in_fast_v = 1;
        p.tw_post = 1;
        p.tw_mul = 1;
        p.sub_tw_stride = m2; // w_m1 = w_m^m2
        p.in_scale_tab = in_tab; p.isc_sv = 1; p.isc_sr = m2;
        e = ntt_run_pass(a, tmp, tw, p, st, batch);
        if (e != cudaSuccess) return e;
        p.in_scale_tab = nullptr;
        // pass 2: rows. vector beta = k1 < m1, element r = j2: in = beta*m2 + r ; out = beta + k*m1
        p.logR = l2;
        p.n_batch = (uint32_t)m1;
        p.n_inner = 1;
        p.T = ntt_pick_T(l2, p.n_batch);
        p.tile_over_batch = 1;
        p.ibs = m2; p.isv = 0; p.isr = 1;
        p.obs = 1; p.osv = 0; p.osr = m1;
        p.in_fast_v = 0;
        p.tw_post = 0;
        p.sub_tw_stride = m1; // w_m2 = w_m^m1
        p.scale_tab = out_tab; p.sc_bs = 1; p.sc_sr = m1;
        e = ntt_run_pass(tmp, a, tw, p, st, batch);
        if (e != cudaSuccess) return e;
        *launches += 2;
    } else {
        // three factors m = m1*m2*m3
        const uint32_t l1 = (log_m + 2) / 3, l2 = (log_m - l1 + 1) / 2, l3 = log_m - l1 - l2;
        const uint64_t m1 = 1ull << l1, m2 = 1ull << l2, m3 = 1ull << l3;
        // pass 1: length-m1 transforms down columns of the m1 x (m2*m3) matrix
        p.logR = l1;
        p.n_batch = 1;
        p.n_inner = (uint32_t)(m2 * m3);
        p.T = ntt_pick_T(l1, p.n_inner);
        p.tile_over_batch = 0;
        p.ibs = 0; p.isv = 1; p.isr = m2 * m3;
        p.obs = 0; p.osv = 1; p.osr = m2 * m3;
        p.in_fast_v = 1;
        p.tw_post = 1;
        p.tw_mul = 1;
        p.sub_tw_stride = m2 * m3;
        p.in_scale_tab = in_tab; p.isc_sv = 1; p.isc_sr = m2 * m3;
        e = ntt_run_pass(a, tmp, tw, p, st, batch);
        if (e != cudaSuccess) return e;
        p.in_scale_tab = nullptr;
        // pass 2: inside each row beta = k1, length-m2 transforms down columns of m2 x m3
        p.logR = l2;
        p.n_batch = (uint32_t)m1;
        p.n_inner = (uint32_t)m3;
        p.T = ntt_pick_T(l2, p.n_inner);
        p.tile_over_batch = 0;
        p.ibs = m2 * m3; p.isv = 1; p.isr = m3;
        p.obs = m2 * m3; p.osv = 1; p.osr = m3;
        p.in_fast_v = 1;
        p.tw_post = 1;
        p.tw_mul = m1; // w_(m2 m3) = w_m^m1
        p.sub_tw_stride = m1 * m3;
        e = ntt_run_pass(tmp, tmp, tw, p, st, batch);
        if (e != cudaSuccess) return e;
        // pass 3: rows (k1, k2) of length m3: in = k1*m2*m3 + k2*m3 + r ; out = k1 + k2*m1 + k*m1*m2
        p.logR = l3;
        p.n_batch = (uint32_t)m1;
        p.n_inner = (uint32_t)m2;
        p.T = ntt_pick_T(l3, p.n_batch);
        p.tile_over_batch = 1;
        p.ibs = m2 * m3; p.isv = m3; p.isr = 1;
        p.obs = 1; p.osv = m1; p.osr = m1 * m2;
        p.in_fast_v = 0;
        p.tw_post = 0;
        p.sub_tw_stride = m1 * m2;
        p.scale_tab = out_tab; p.sc_bs = 1; p.sc_sv = m1; p.sc_sr = m1 * m2;
        e = ntt_run_pass(tmp, a, tw, p, st, batch);
        if (e != cudaSuccess) return e;
        *launches += 3;
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------ Lagrange basis at a point (setup generator)
// u[i] = L_i(t) for the domain of m = 2^log_m roots of unity: libfqfft's
// _basic_radix2_evaluate_all_lagrange_polynomials (basic_radix2_domain_aux.tcc:186-239),
//     u[i] = Z(t) / m * w^i / (t - w^i),   Z(t) = t^m - 1,
// which the reference computes with m serial field inversions (r1cs_to_qap_instance_map_with_evaluation,
// called once per trusted setup, r1cs_gg_ppzksnark.tcc:243-247). Here: one thread per LAGRANGE_CHUNK
// consecutive i, the chunk's denominators inverted together (Montgomery's trick, one Fermat inversion).
// seed[0] = Z(t)/m, seed[1] = t, flag = 1 if t^m == 1 (t is a domain point: the caller falls back).
static constexpr uint32_t LAGRANGE_CHUNK = 32;

__global__ void k_lagrange_seed(Fr t, uint32_t log_m, Fr *seed, int *flag)
{
    uint32_t x[12], one[12], mm[12], two[12];
    fr_load(x, &t);
    fr_set_const(one, FR_ONE);
    for (uint32_t i = 0; i < log_m; i++) fr_sqr_regs(x, x); // t^m
    fr_sub_regs(x, x, one);
    uint32_t nz = 0;
#pragma unroll
    for (int k = 0; k < 12; k++) nz |= x[k];
    *flag = nz == 0;
    fr_add_regs(two, one, one);
    fr_set_const(mm, FR_ONE);
    for (uint32_t i = 0; i < log_m; i++) fr_mul_regs(mm, mm, two);
    fr_inv(mm, mm);
    fr_mul_regs(x, x, mm);
    fr_store(seed, x);
    fr_store(seed + 1, t.l);
}

__global__ void __launch_bounds__(128) k_lagrange(const Fr *__restrict__ tw, uint32_t log_m, const Fr *seed, Fr *out)
{
    const size_t m = (size_t)1 << log_m;
    const size_t i0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * LAGRANGE_CHUNK;
    if (i0 >= m) return;
    const uint32_t cnt = (uint32_t)(m - i0 < LAGRANGE_CHUNK ? m - i0 : LAGRANGE_CHUNK);
    uint32_t t[12], acc[12], w[12], d[12];
    fr_load(t, seed + 1);
    fr_set_const(acc, FR_ONE);
    for (uint32_t j = 0; j < cnt; j++) { // forward: out[i] = product of the denominators before i
        fr_store(out + i0 + j, acc);
        ntt_twiddle(w, tw, i0 + j, log_m, 0);
        fr_sub_regs(d, t, w);
        fr_mul_regs(acc, acc, d);
    }
    fr_inv(acc, acc); // != 0: t is not a domain point
    uint32_t l0[12];
    fr_load(l0, seed);
    for (uint32_t j = cnt; j-- > 0;) { // backward
        uint32_t pre[12], u[12];
        fr_load(pre, out + i0 + j);
        fr_mul_regs(u, acc, pre); // 1 / (t - w^i)
        ntt_twiddle(w, tw, i0 + j, log_m, 0);
        fr_sub_regs(d, t, w);
        fr_mul_regs(acc, acc, d);
        fr_mul_regs(u, u, w);
        fr_mul_regs(u, u, l0);
        fr_store(out + i0 + j, u);
    }
}

// ------------------------------------------------------------------ distributed 4-step
// One transform of m = m1*m2 elements over `world` ranks (a power of two dividing m1 and m2) with
// ONE exchange. Layouts of the distributed vector (c1 = m1/world, c2 = m2/world):
//   layout A: rank g holds the elements whose index j = j1*m2 + j2 has j2 in [g*c2, (g+1)*c2),
//             stored [j1][j2 - g*c2]                      (m1 x c2, "column block")
//   layout B: rank h holds the elements whose index k = k1 + m1*k2 has k1 in [h*c1, (h+1)*c1),
//             stored [k1 - h*c1][k2]                      (c1 x m2)
// forward (FFT, cosetFFT):   layout A -> layout B;   inverse (iFFT, icosetFFT): layout B -> layout A.
// phase 0 = local sub-transforms + inter-step twiddle, output written as `world` contiguous chunks
// (chunk d goes to rank d); then the caller exchanges chunks (all-to-all: chunk s of the receive
// buffer comes from rank s); phase 1 = local sub-transforms on the received chunks.
// Coset multiplication (forward) and the 1/m, g^-i scaling (inverse) are fused into the first /
// last pass. `in` and `out` are different buffers of m/world elements.
inline cudaError_t ntt_dist_phase(NttTables &tabs, const Fr *in, Fr *out, uint32_t log_m, uint32_t flags, const Fr *coset_g,
                                  uint32_t world, uint32_t rank, uint32_t phase, cudaStream_t st, unsigned *launches)
{
    const uint32_t inverse = flags & 1u, coset = (flags >> 1) & 1u;
    uint32_t lg = 0;
    while ((1u << lg) < world) lg++;
    const uint32_t l1 = (log_m + 1) / 2, l2 = log_m - l1;
    if ((1u << lg) != world || lg >= l2 || l1 > 12 || rank >= world || phase > 1) return cudaErrorInvalidValue;
    const uint64_t m1 = 1ull << l1, m2 = 1ull << l2, c1 = m1 >> lg, c2 = m2 >> lg;
    const Fr *tw = nullptr;
    cudaError_t e = ntt_get_twiddles(tabs, log_m, st, &tw, launches);
    if (e != cudaSuccess) return e;
    const Fr *ctab = nullptr;
    Fr one_g;
    memcpy(one_g.l, FR_ONE_HOST, sizeof(Fr));
    if (coset || inverse) {
        e = ntt_get_coset(tabs, log_m, coset ? *coset_g : one_g, inverse, st, &ctab, launches);
        if (e != cudaSuccess) return e;
    }
    NttPass p{};
    p.log_m = log_m;
    p.inverse = inverse;
    if (!inverse && phase == 0) {
        // columns of layout A: vector u = local column, element r = j1
        p.logR = l1; p.n_batch = 1; p.n_inner = (uint32_t)c2; p.T = ntt_pick_T(l1, p.n_inner);
        p.tile_over_batch = 0; p.in_fast_v = 1;
        p.isv = 1; p.isr = c2;
        p.osv = 1; p.osr = c2;                       // [k1][j2 local]: chunk d = rows k1 in [d*c1, (d+1)*c1)
        p.tw_post = 1; p.tw_mul = 1; p.tw_vec_off = (uint64_t)rank * c2;
        p.sub_tw_stride = m2;
        if (coset) { p.in_scale_tab = ctab; p.isc_sv = 1; p.isc_sr = m2; p.isc_off = (uint64_t)rank * c2; }
    } else if (!inverse) {
        // rows of layout B from the received chunks [s][k1 local][j2 local]: vector beta = k1 local, element r = j2
        p.logR = l2; p.n_batch = (uint32_t)c1; p.n_inner = 1; p.T = ntt_pick_T(l2, p.n_batch);
        p.tile_over_batch = 1; p.in_fast_v = 0;
        p.ibs = c2; p.isr = 1; p.r_lo_bits = l2 - lg; p.isr_hi = c1 * c2;
        p.obs = m2; p.osr = 1;
        p.sub_tw_stride = m1;
    } else if (phase == 0) {
        // rows of layout B: vector beta = k1 local, element r = k2; twiddle w^-(k1 * j2)
        p.logR = l2; p.n_batch = (uint32_t)c1; p.n_inner = 1; p.T = ntt_pick_T(l2, p.n_batch);
        p.tile_over_batch = 1; p.in_fast_v = 0;
        p.ibs = m2; p.isr = 1;
        p.obs = c2; p.osr = 1; p.k_lo_bits = l2 - lg; p.osr_hi = c1 * c2; // chunk d = columns j2 in [d*c2, (d+1)*c2)
        p.tw_post = 1; p.tw_mul = 1; p.tw_use_beta = 1; p.tw_vec_off = (uint64_t)rank * c1;
        p.sub_tw_stride = m1;
    } else {
        // columns of layout A from the received chunks [s][k1 local][j2 local]: vector u = j2 local, element r = k1
        p.logR = l1; p.n_batch = 1; p.n_inner = (uint32_t)c2; p.T = ntt_pick_T(l1, p.n_inner);
        p.tile_over_batch = 0; p.in_fast_v = 1;
        p.isv = 1; p.isr = c2; p.r_lo_bits = l1 - lg; p.isr_hi = c1 * c2;
        p.osv = 1; p.osr = c2;
        p.sub_tw_stride = m2;
        p.scale_tab = ctab; p.sc_sv = 1; p.sc_sr = m2; p.sc_off = (uint64_t)rank * c2; // 1/m (and g^-j)
    }
    e = ntt_run_pass(in, out, tw, p, st);
    (*launches)++;
    return e;
}

} // namespace zecale
