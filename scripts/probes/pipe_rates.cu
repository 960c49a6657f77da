// Issue-rate probe for the next round's multiplier choice: dependent chains of DFMA (FP64 pipe) and of
// IMAD.WIDE.U32 (integer multiply pipe) with 8 independent accumulators per thread, alone and mixed in the
// same warp. Prints operations per second over the whole GPU. Build: nvcc -gencode arch=compute_100a,code=sm_100a
// -O3 -o pipe_rates pipe_rates.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;

__global__ void k_dfma(double *out, double a, double b)
{
    double x[8];
    for (int k = 0; k < 8; k++) x[k] = threadIdx.x + k;
    for (int i = 0; i < ITERS; i++)
#pragma unroll
        for (int k = 0; k < 8; k++) x[k] = fma(x[k], a, b);
    double s = 0;
    for (int k = 0; k < 8; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_imad(uint64_t *out, uint32_t a, uint32_t b)
{
    uint64_t x[8];
    for (int k = 0; k < 8; k++) x[k] = threadIdx.x + k;
    for (int i = 0; i < ITERS; i++)
#pragma unroll
        for (int k = 0; k < 8; k++) x[k] = (uint64_t)((uint32_t)x[k] ^ a) * b + x[k]; // IMAD.WIDE.U32 + LOP3
    uint64_t s = 0;
    for (int k = 0; k < 8; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_mixed(uint64_t *out, uint32_t a, uint32_t b, double da, double db)
{
    uint64_t x[4];
    double y[8];
    for (int k = 0; k < 4; k++) x[k] = threadIdx.x + k;
    for (int k = 0; k < 8; k++) y[k] = threadIdx.x + k;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int k = 0; k < 4; k++) x[k] = (uint64_t)((uint32_t)x[k] ^ a) * b + x[k];
#pragma unroll
        for (int k = 0; k < 8; k++) y[k] = fma(y[k], da, db);
    }
    uint64_t s = 0;
    for (int k = 0; k < 4; k++) s += x[k];
    for (int k = 0; k < 8; k++) s += (uint64_t)y[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F> static float time_ms(F f)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        f();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, tpb = 256;
    void *buf;
    cudaMalloc(&buf, (size_t)sms * 8 * tpb * 8);
    for (int bps = 1; bps <= 8; bps *= 2) {
        const int blocks = sms * bps;
        const double n = (double)blocks * tpb * ITERS;
        float t1 = time_ms([&] { k_dfma<<<blocks, tpb>>>((double *)buf, 1.0000001, 1e-9); });
        float t2 = time_ms([&] { k_imad<<<blocks, tpb>>>((uint64_t *)buf, 0x9e3779b9u, 0x85ebca6bu); });
        float t3 = time_ms([&] { k_mixed<<<blocks, tpb>>>((uint64_t *)buf, 0x9e3779b9u, 0x85ebca6bu, 1.0000001, 1e-9); });
        printf("{\"sms\": %d, \"warps_per_sm\": %d, \"dfma_per_s\": %.4g, \"imad_wide_per_s\": %.4g, "
               "\"mixed_imad_wide_per_s\": %.4g, \"mixed_dfma_per_s\": %.4g}\n",
               sms, bps * tpb / 32, n * 8 / (t1 * 1e-3), n * 8 / (t2 * 1e-3), n * 4 / (t3 * 1e-3), n * 8 / (t3 * 1e-3));
    }
    return 0;
}
