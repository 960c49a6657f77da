// Product-only prototype of a BW6-761 Fq Montgomery multiplication on the FP64 pipe (DESIGN.md "what comes
// next", lead 1): 15 limbs of 52 bits held in doubles, every limb product split exactly by two
// round-toward-zero DFMAs, column sums kept as 64-bit integer sums of the raw double bit patterns,
// word-serial Montgomery reduction in product-scanning order (R = 2^780). The integer model and the
// exactness checker are scripts/probes/dfma_model.py. NOT part of the product: nothing in zecale_b200/ uses it.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o dfma_mul dfma_mul.cu
// Run:   ./dfma_mul dump.bin [iters]      writes the checker's dump, prints one JSON line per timing
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#include "dfma_consts.h"
#define DFMA_N 15
#define DFMA_Q DFMA52_Q
#define DFMA_Q_HOST DFMA52_Q_HOST
#define DFMA_ONE_HOST DFMA52_ONE_HOST
#define DFMA_QINV DFMA52_QINV

typedef unsigned long long u64;
constexpr int N = DFMA_N;
constexpr u64 EXP_HI = (u64)(1023 + 104) << 52; // bit pattern of 2^104
constexpr u64 EXP_LO = (u64)(1023 + 52) << 52;  // bit pattern of 2^52
constexpr u64 MASK52 = ((u64)1 << 52) - 1;

// cur += low 52 bits of x*y, nxt += high 52 bits, both as raw bit patterns (exponent fields taken out per column)
__device__ __forceinline__ void mac(double x, double y, u64 &cur, u64 &nxt)
{
    const double hi = __fma_rz(x, y, 0x1p104);
    const double sub = 0x1.0000000000001p104 - hi; // 2^104 + 2^52 - hi: exact
    const double lo = __fma_rz(x, y, sub);
    nxt += (u64)__double_as_longlong(hi);
    cur += (u64)__double_as_longlong(lo);
}

__device__ __forceinline__ double limb_to_double(u64 t) // t < 2^52
{
    return __longlong_as_double((long long)(t | EXP_LO)) - 0x1p52;
}

// r = a * b * 2^-780 mod q, r < q + q / 2^16 for a, b < 2 q; limbs of r below 2^52
__device__ __forceinline__ void fq52_mul(double (&r)[N], const double (&a)[N], const double (&b)[N])
{
    double m[N];
    u64 cur = 0, carry = 0;
#pragma unroll
    for (int k = 0; k < 2 * N; k++) {
        u64 nxt = 0;
        int n = 0;
#pragma unroll
        for (int i = 0; i < N; i++) {
            const int j = k - i;
            if (j >= 0 && j < N) {
                mac(a[i], b[j], cur, nxt);
                n++;
            }
        }
#pragma unroll
        for (int i = 0; i < N; i++) {
            const int j = k - i;
            if (i < k && j >= 0 && j < N) {
                mac(m[i], DFMA_Q[j], cur, nxt);
                n++;
            }
        }
        if (k < N) {
            u64 col = cur + (carry - (u64)n * EXP_LO);
            const double t = limb_to_double(col & MASK52);
            // m_k = t * (-1/q) mod 2^52
            const double hi = __fma_rz(t, DFMA_QINV, 0x1p104);
            const double lo = __fma_rz(t, DFMA_QINV, 0x1.0000000000001p104 - hi);
            m[k] = lo - 0x1p52;
            mac(m[k], DFMA_Q[0], col, nxt);
            col -= EXP_LO;                 // low 52 bits are zero now
            carry = col >> 52;
            cur = nxt - (u64)(n + 1) * EXP_HI;
        } else {
            const u64 col = cur + (carry - (u64)n * EXP_LO);
            r[k - N] = limb_to_double(col & MASK52);
            carry = col >> 52;
            cur = nxt - (u64)n * EXP_HI;
        }
    }
}

__global__ void __launch_bounds__(128) k_check(const u64 *in, u64 *out, int n, int iters)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double a[N], b[N], r[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
        a[i] = (double)in[(size_t)t * 2 * N + i];
        b[i] = (double)in[(size_t)t * 2 * N + N + i];
    }
    for (int it = 0; it < iters; it++) {
        fq52_mul(r, a, b);
#pragma unroll
        for (int i = 0; i < N; i++) a[i] = r[i];
    }
#pragma unroll
    for (int i = 0; i < N; i++) out[(size_t)t * N + i] = (u64)a[i];
}

template <int MINB>
__global__ void __launch_bounds__(128, MINB) k_chain(double *sink, int iters)
{
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    double a[N], b[N], r[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
        a[i] = (double)((t * 2654435761u >> (i & 7)) + i);
        b[i] = (double)(t + 3 * i + 1);
    }
    for (int it = 0; it < iters; it++) {
        fq52_mul(r, a, b);
#pragma unroll
        for (int i = 0; i < N; i++) a[i] = r[i];
    }
    if (a[0] == 12345.0 && a[5] == 77.0)
        for (int i = 0; i < N; i++) sink[i] = a[i]; // keep the chain alive
}

static u64 rng_state = 0x9E3779B97F4A7C15ull;
static u64 rnd()
{
    rng_state ^= rng_state << 13;
    rng_state ^= rng_state >> 7;
    rng_state ^= rng_state << 17;
    return rng_state;
}

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e_ = (x);                                                      \
        if (e_ != cudaSuccess) {                                                   \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));               \
            return 2;                                                              \
        }                                                                          \
    } while (0)

template <int MINB> static int time_chain(double *sink, int blocks, int iters)
{
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    k_chain<MINB><<<blocks, 128>>>(sink, 8);
    CK(cudaEventRecord(e0));
    k_chain<MINB><<<blocks, 128>>>(sink, iters);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double prods = (double)blocks * 128 * iters;
    printf("{\"probe\": \"dfma_mul chain\", \"min_blocks_per_sm\": %d, \"blocks\": %d, \"threads\": 128, \"iters\": %d, \"ms\": %.4f, "
           "\"fq_products_per_s\": %.4e, \"equiv_32x32_products_per_s\": %.4e}\n",
           MINB, blocks, iters, ms, prods / (ms * 1e-3), prods * 1176.0 / (ms * 1e-3));
    return 0;
}

int main(int argc, char **argv)
{
    const char *dump = argc > 1 ? argv[1] : "dfma_dump.bin";
    const int n = 4096;
    FILE *f = fopen(dump, "wb");
    if (!f) return 1;
    for (int iters : {1, 40}) {
        std::vector<u64> in((size_t)n * 2 * N), out((size_t)n * N);
        for (int t = 0; t < n; t++)
            for (int h = 0; h < 2; h++) {
                u64 *v = &in[(size_t)t * 2 * N + h * N];
                for (int i = 0; i < N; i++) v[i] = rnd() & MASK52;
                v[N - 1] = rnd() % (2 * DFMA_Q_HOST[N - 1]); // value < 2 q (+ 2^728)
                if (t < 4) // edge cases: zero, one (Montgomery form), q - 1 style all-ones limbs
                    for (int i = 0; i < N; i++) v[i] = t == 0 ? 0 : t == 1 ? DFMA_ONE_HOST[i] : t == 2 ? DFMA_Q_HOST[i] - (i == 0) : (i < N - 1 ? MASK52 : DFMA_Q_HOST[N - 1]);
            }
        u64 *d_in, *d_out;
        CK(cudaMalloc(&d_in, in.size() * 8));
        CK(cudaMalloc(&d_out, out.size() * 8));
        CK(cudaMemcpy(d_in, in.data(), in.size() * 8, cudaMemcpyHostToDevice));
        k_check<<<(n + 127) / 128, 128>>>(d_in, d_out, n, iters);
        CK(cudaMemcpy(out.data(), d_out, out.size() * 8, cudaMemcpyDeviceToHost));
        const u64 hdr[4] = {(u64)n, (u64)iters, 52, (u64)N};
        fwrite(hdr, 8, 4, f);
        for (int t = 0; t < n; t++) {
            fwrite(&in[(size_t)t * 2 * N], 8, 2 * N, f);
            fwrite(&out[(size_t)t * N], 8, N, f);
        }
        CK(cudaFree(d_in));
        CK(cudaFree(d_out));
    }
    fclose(f);
    double *sink;
    CK(cudaMalloc(&sink, 8 * N));
    const int iters = argc > 2 ? atoi(argv[2]) : 2000;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    for (int mult : {1, 2, 3, 4, 6}) {
        if (time_chain<2>(sink, sms * mult * 2, iters)) return 2;
        if (time_chain<3>(sink, sms * mult * 3, iters)) return 2;
        if (time_chain<4>(sink, sms * mult * 4, iters)) return 2;
    }
    return 0;
}
