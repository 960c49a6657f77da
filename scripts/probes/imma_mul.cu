// Product-only prototype: BW6-761 Fq Montgomery multiplication with the two constant products of the reduction
// (m = P_lo * q' mod R, U = m * q) on the tensor cores. A warp holds 32 multiplications, one per lane; the lane computes
// P = a * b with the integer pipe (576 IMAD.WIDE), the warp multiplies the 32 x 96 byte matrix of the P_lo (then m) limbs
// by constant Toeplitz matrices with mma.sync.m16n8k32.u8 (s32 accumulators), column sums go back to their lane through
// shared memory as radix-2^16 digits and are carry-normalised there. Integer model, tables and checker:
// scripts/probes/imma_model.py. NOT part of the product: nothing in zecale_b200/ uses it.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o imma_mul imma_mul.cu
// Run:   ./imma_mul dump.bin [iters]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#include "imma_consts.h"

constexpr int N = 24;
constexpr int A_STRIDE = 28;  // words per row of the operand staging (112 B: conflict-free STS.128 and fragment loads)
constexpr int D_STRIDE = 52;  // words per row of the digit staging (48 digits = 12 column tiles + pad; conflict-free both ways)
#ifndef DEPTH
#define DEPTH 2
#endif
#define IMMA_PRODUCT_(d) imma_full_product_d##d
#define IMMA_PRODUCT(d) IMMA_PRODUCT_(d)
constexpr int WARPS = 4;

struct WarpStage {
    uint32_t a[32 * A_STRIDE];
    uint32_t d[32 * D_STRIDE];
};

__device__ __forceinline__ void mma_u8(int (&acc)[4], const uint32_t (&a)[4], uint2 b)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

// digits of the 32 rows' products with a constant, column tiles [NT0, NT0 + 12): W.d[row][4 (nt - NT0) + q]
template <int NT0>
__device__ __forceinline__ void toeplitz_gemm(WarpStage &W, const uint2 *__restrict__ tiles, int lane)
{
    const int g = lane >> 2, q = lane & 3;
    uint32_t A[2][3][4];
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int ks = 0; ks < 3; ks++) {
            const uint32_t *r0 = W.a + (16 * mt + g) * A_STRIDE + 8 * ks + q;
            A[mt][ks][0] = r0[0];
            A[mt][ks][1] = r0[8 * A_STRIDE];
            A[mt][ks][2] = r0[4];
            A[mt][ks][3] = r0[8 * A_STRIDE + 4];
        }
#pragma unroll
    for (int nt = NT0; nt < NT0 + 12; nt++) {
        int acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
#pragma unroll
        for (int ks = 0; ks < 3; ks++) {
            const int d = nt - 4 * ks;
            if (d >= 0 && d < 16) {
                const uint2 b = tiles[d * 32 + lane];
                mma_u8(acc[0], A[0][ks], b);
                mma_u8(acc[1], A[1][ks], b);
            }
        }
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            W.d[(16 * mt + g) * D_STRIDE + 4 * (nt - NT0) + q] = (uint32_t)acc[mt][0] + ((uint32_t)acc[mt][1] << 8);
            W.d[(16 * mt + g + 8) * D_STRIDE + 4 * (nt - NT0) + q] = (uint32_t)acc[mt][2] + ((uint32_t)acc[mt][3] << 8);
        }
    }
}

// limbs = the lane's 48 staged digits (+ carry in); returns the carry out
__device__ __forceinline__ uint32_t normalise(uint32_t (&limbs)[24], const uint32_t *digits, uint32_t carry_in = 0)
{
    uint32_t d[48];
#pragma unroll
    for (int k4 = 0; k4 < 12; k4++) {
        const uint4 v = reinterpret_cast<const uint4 *>(digits)[k4];
        d[4 * k4] = v.x;
        d[4 * k4 + 1] = v.y;
        d[4 * k4 + 2] = v.z;
        d[4 * k4 + 3] = v.w;
    }
    return imma_normalise24(limbs, d, carry_in);
}

__device__ __forceinline__ void stage_operand(WarpStage &W, const uint32_t *x, int lane)
{
    uint4 *row = reinterpret_cast<uint4 *>(W.a + lane * A_STRIDE);
#pragma unroll
    for (int k = 0; k < 6; k++) row[k] = make_uint4(x[4 * k], x[4 * k + 1], x[4 * k + 2], x[4 * k + 3]);
}

// r = a * b / 2^768 mod q, canonical; all 32 lanes of the warp must call it together
__device__ __forceinline__ void fq_mul_imma(uint32_t (&r)[N], const uint32_t (&a)[N], const uint32_t (&b)[N], WarpStage &W,
                                            const uint2 *tq, const uint2 *tn, const uint32_t *qmod, int lane, uint32_t zero)
{
    uint32_t P[2 * N];
    IMMA_PRODUCT(DEPTH)(P, a, b, zero);
    uint32_t nz = 0;
#pragma unroll
    for (int k = 0; k < N; k++) nz |= P[k];
    stage_operand(W, P, lane);
    __syncwarp();
    toeplitz_gemm<0>(W, tq, lane);
    __syncwarp();
    uint32_t m[N];
    normalise(m, W.d + lane * D_STRIDE);
    stage_operand(W, m, lane);
    __syncwarp();
    toeplitz_gemm<0>(W, tn, lane);
    __syncwarp();
    uint32_t lo[N], hi[N];
    const uint32_t c = normalise(lo, W.d + lane * D_STRIDE);   // low half of m q: only its carry is used
    __syncwarp();
    toeplitz_gemm<12>(W, tn, lane);
    __syncwarp();
    normalise(hi, W.d + lane * D_STRIDE, c);
    __syncwarp();
    // T = P_hi + U_hi + (P_lo != 0)
    unsigned long long carry = nz ? 1u : 0u;
    uint32_t T[N];
#pragma unroll
    for (int k = 0; k < N; k++) {
        const unsigned long long t = (unsigned long long)P[N + k] + hi[k] + carry;
        T[k] = (uint32_t)t;
        carry = t >> 32;
    }
    // conditional subtract
    uint32_t D[N];
    long long borrow = 0;
#pragma unroll
    for (int k = 0; k < N; k++) {
        const long long t = (long long)T[k] - qmod[k] + borrow;
        D[k] = (uint32_t)t;
        borrow = t >> 32;
    }
#pragma unroll
    for (int k = 0; k < N; k++) r[k] = borrow ? T[k] : D[k];
}

__constant__ uint32_t Q_DEV[N];

__global__ void __launch_bounds__(32 * WARPS) k_check(const uint32_t *in, uint32_t *out, const uint2 *tq_g, const uint2 *tn_g, int iters, uint32_t zero)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint2 *tq = reinterpret_cast<uint2 *>(smem), *tn = tq + 16 * 32;
    WarpStage *stage = reinterpret_cast<WarpStage *>(tn + 16 * 32);
    for (int i = threadIdx.x; i < 16 * 32; i += blockDim.x) {
        tq[i] = tq_g[i];
        tn[i] = tn_g[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a[N], b[N], r[N];
#pragma unroll
    for (int k = 0; k < N; k++) {
        a[k] = in[(size_t)t * 2 * N + k];
        b[k] = in[(size_t)t * 2 * N + N + k];
    }
    for (int it = 0; it < iters; it++) {
        fq_mul_imma(r, a, b, stage[threadIdx.x >> 5], tq, tn, Q_DEV, lane, zero);
#pragma unroll
        for (int k = 0; k < N; k++) a[k] = r[k];
    }
#pragma unroll
    for (int k = 0; k < N; k++) out[(size_t)t * N + k] = a[k];
}

template <int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_chain(uint32_t *sink, const uint2 *tq_g, const uint2 *tn_g, int iters, uint32_t zero)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint2 *tq = reinterpret_cast<uint2 *>(smem), *tn = tq + 16 * 32;
    WarpStage *stage = reinterpret_cast<WarpStage *>(tn + 16 * 32);
    for (int i = threadIdx.x; i < 16 * 32; i += blockDim.x) {
        tq[i] = tq_g[i];
        tn[i] = tn_g[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a[N], b[N], r[N];
#pragma unroll
    for (int k = 0; k < N; k++) {
        a[k] = (t * 2654435761u >> (k & 7)) + k;
        b[k] = t + 3 * k + 1;
    }
    a[N - 1] &= 0xffffff;
    b[N - 1] &= 0xffffff;
    for (int it = 0; it < iters; it++) {
        fq_mul_imma(r, a, b, stage[threadIdx.x >> 5], tq, tn, Q_DEV, lane, zero);
#pragma unroll
        for (int k = 0; k < N; k++) a[k] = r[k];
    }
    if (a[0] == 0x12345678u && a[5] == 77u)
        for (int k = 0; k < N; k++) sink[k] = a[k];
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd()
{
    rng_state ^= rng_state << 13;
    rng_state ^= rng_state >> 7;
    rng_state ^= rng_state << 17;
    return rng_state;
}

#define CK(x)                                                        \
    do {                                                             \
        cudaError_t e_ = (x);                                        \
        if (e_ != cudaSuccess) {                                     \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); \
            return 2;                                                \
        }                                                            \
    } while (0)

constexpr size_t SMEM = 2 * 16 * 32 * sizeof(uint2) + WARPS * sizeof(WarpStage);

template <int MINB> static int time_chain(uint32_t *sink, const uint2 *tq, const uint2 *tn, int blocks, int iters)
{
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaFuncSetAttribute(k_chain<MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    k_chain<MINB><<<blocks, 32 * WARPS, SMEM>>>(sink, tq, tn, 8, 0u);
    CK(cudaEventRecord(e0));
    k_chain<MINB><<<blocks, 32 * WARPS, SMEM>>>(sink, tq, tn, iters, 0u);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double prods = (double)blocks * 32 * WARPS * iters;
    printf("{\"probe\": \"imma_mul chain\", \"rows_in_flight\": %d, \"min_blocks_per_sm\": %d, \"blocks\": %d, \"threads\": %d, \"iters\": %d, \"ms\": %.4f, "
           "\"fq_products_per_s\": %.4e, \"equiv_32x32_products_per_s\": %.4e}\n",
           DEPTH, MINB, blocks, 32 * WARPS, iters, ms, prods / (ms * 1e-3), prods * 1176.0 / (ms * 1e-3));
    return 0;
}

int main(int argc, char **argv)
{
    const char *dump = argc > 1 ? argv[1] : "imma_dump.bin";
    const int n = 4096;
    uint2 *tq, *tn;
    CK(cudaMalloc(&tq, sizeof(IMMA_TQ_HOST)));
    CK(cudaMalloc(&tn, sizeof(IMMA_TN_HOST)));
    CK(cudaMemcpy(tq, IMMA_TQ_HOST, sizeof(IMMA_TQ_HOST), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(tn, IMMA_TN_HOST, sizeof(IMMA_TN_HOST), cudaMemcpyHostToDevice));
    CK(cudaMemcpyToSymbol(Q_DEV, IMMA_Q_HOST, sizeof(IMMA_Q_HOST)));
    CK(cudaFuncSetAttribute(k_check, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    FILE *f = fopen(dump, "wb");
    if (!f) return 1;
    for (int iters : {1, 40}) {
        std::vector<uint32_t> in((size_t)n * 2 * N), out((size_t)n * N);
        for (int t = 0; t < n; t++)
            for (int h = 0; h < 2; h++) {
                uint32_t *v = &in[(size_t)t * 2 * N + h * N];
                for (int i = 0; i < N; i++) v[i] = (uint32_t)rnd();
                v[N - 1] = (uint32_t)(rnd() % IMMA_Q_HOST[N - 1]); // value < q
                if (t < 4) // edge cases: zero, one (Montgomery form), q - 1, all-ones low limbs
                    for (int i = 0; i < N; i++)
                        v[i] = t == 0 ? 0 : t == 1 ? IMMA_ONE_HOST[i] : t == 2 ? IMMA_Q_HOST[i] - (i == 0) : (i < N - 1 ? 0xffffffffu : IMMA_Q_HOST[N - 1] - 1);
            }
        uint32_t *d_in, *d_out;
        CK(cudaMalloc(&d_in, in.size() * 4));
        CK(cudaMalloc(&d_out, out.size() * 4));
        CK(cudaMemcpy(d_in, in.data(), in.size() * 4, cudaMemcpyHostToDevice));
        k_check<<<n / (32 * WARPS), 32 * WARPS, SMEM>>>(d_in, d_out, tq, tn, iters, 0u);
        CK(cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost));
        const uint64_t hdr[2] = {(uint64_t)n, (uint64_t)iters};
        fwrite(hdr, 8, 2, f);
        for (int t = 0; t < n; t++) {
            fwrite(&in[(size_t)t * 2 * N], 4, 2 * N, f);
            fwrite(&out[(size_t)t * N], 4, N, f);
        }
        CK(cudaFree(d_in));
        CK(cudaFree(d_out));
    }
    fclose(f);
    uint32_t *sink;
    CK(cudaMalloc(&sink, 4 * N));
    const int iters = argc > 2 ? atoi(argv[2]) : 1000;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    for (int mult : {1, 2, 4}) {
        if (time_chain<2>(sink, tq, tn, sms * mult * 2, iters)) return 2;
        if (time_chain<3>(sink, tq, tn, sms * mult * 3, iters)) return 2;
        if (time_chain<4>(sink, tq, tn, sms * mult * 4, iters)) return 2;
    }
    return 0;
}
