// Micro-benchmark: issue efficiency of the Frequency encoding alone (no TMEM / MMA), one query per thread.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../neural-radiance-caching_b200/csrc encode_ipc.cu -o encode_ipc
#include <cstdio>
#include <cuda_runtime.h>
#include "nrc_encode.cuh"
using namespace nrcb;

template <int MODE>
__global__ void __launch_bounds__(512) k_encode(const float* __restrict__ in, uint32_t* __restrict__ out, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float q[15];
#pragma unroll
        for (int j = 0; j < 15; ++j) q[j] = in[(size_t)i * 15 + j];
        uint32_t acc = 0;
        if (MODE == 0) {          // full encoding
            uint32_t e[40];
            encode_frequency(q, e);
#pragma unroll
            for (int j = 0; j < 40; ++j) acc ^= e[j];
        } else if (MODE == 1) {   // tri-wave only
#pragma unroll
            for (int d = 0; d < 3; ++d)
#pragma unroll
                for (int f = 0; f < 12; f += 2) acc ^= pack_h2(triwave(q[d], f), triwave(q[d], f + 1));
        } else {                  // one-blob only
#pragma unroll
            for (int d = 0; d < 6; ++d) {
                float y[4];
                oneblob4<true>(q[3 + d], y);
                acc ^= pack_h2(y[0], y[1]) ^ pack_h2(y[2], y[3]);
            }
        }
        out[i] = acc;
    }
}

int main() {
    const uint32_t n = 3744000;
    float* in; uint32_t* out;
    cudaMalloc(&in, (size_t)n * 60); cudaMalloc(&out, (size_t)n * 4);
    cudaMemset(in, 0x3c, (size_t)n * 60);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int mode = 0; mode < 3; ++mode)
        for (int ctas = 1; ctas <= 4; ctas *= 2) {
            float best = 1e9;
            for (int r = 0; r < 5; ++r) {
                cudaEventRecord(a);
                if (mode == 0) k_encode<0><<<148 * ctas, 512>>>(in, out, n);
                if (mode == 1) k_encode<1><<<148 * ctas, 512>>>(in, out, n);
                if (mode == 2) k_encode<2><<<148 * ctas, 512>>>(in, out, n);
                cudaEventRecord(b); cudaEventSynchronize(b);
                float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
            }
            printf("mode %d (0 full, 1 tri-wave, 2 one-blob)  %d CTAs x 512 thr per SM: %.4f ms  err=%d\n", mode, ctas, best, (int)cudaGetLastError());
        }
    return 0;
}
