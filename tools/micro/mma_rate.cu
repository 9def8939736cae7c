// Micro-benchmark: tcgen05.mma issue rate / dependent-chain latency (M=128, K=16, kind::f16, fp32 accumulate).
//   NACC  independent accumulators used round-robin by each issuing warp (1 = one dependent chain)
//   WARPS warps issuing concurrently (each to its own accumulators)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../neural-radiance-caching_b200/csrc mma_rate.cu -o mma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "sm100_prims.cuh"
using namespace nrcb;

template <int N, int NACC, int WARPS, bool TS>
__global__ void __launch_bounds__(128, 1) k_rate(long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 16384; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u;
    if (tid == 0) { for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(&slot, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = slot;
    constexpr int REPS = 256;
    constexpr uint32_t idesc = umma_idesc_f16(128, N, 0, 0);
    const uint32_t sb = smem_u32(smem);
    const uint64_t ad = umma_desc(sb, 2048, 128), bd = umma_desc(sb + 32768, N * 16, 128);
    if (warp < WARPS) {
        const bool leader = elect_one();
        const uint32_t d0 = tb + warp * (N <= 64 ? 64 : 128) * 1;   // accumulators of this warp: d0 + i * N (columns)
        __syncwarp();
        long long t0 = clock64(), t1 = 0;
        if (leader) {
#pragma unroll 8
            for (int r = 0; r < REPS; ++r) {
                const uint32_t d = d0 + (NACC == 1 ? 0 : (r % NACC) * (512 / (WARPS * NACC) >= N ? N : 0));
                if (TS) umma_ts(d, tb + 448, bd, idesc, 1); else umma_ss(d, ad, bd, idesc, 1);
            }
            t1 = clock64();
            umma_commit(&bar[warp]);
        }
        __syncwarp();
        mbar_wait(&bar[warp], 0);
        long long t2 = clock64();
        if (leader) { out[warp * 2] = t1 - t0; out[warp * 2 + 1] = t2 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 512);
}

template <int N, int NACC, int WARPS, bool TS>
void run(const char* what) {
    long long* d; cudaMalloc(&d, 64); cudaMemset(d, 0, 64);
    cudaFuncSetAttribute(k_rate<N, NACC, WARPS, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    for (int i = 0; i < 2; ++i) k_rate<N, NACC, WARPS, TS><<<1, 128, 65536>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[8]; cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
    printf("%-34s N=%3d nacc=%d warps=%d %s: issue %6.1f cyc/MMA, complete %6.1f cyc/MMA (warp 0)%s\n", what, N, NACC, WARPS, TS ? "TS" : "SS",
           h[0] / 256.0, h[1] / 256.0, e == cudaSuccess ? "" : "  CUDA ERROR");
    if (WARPS > 1) printf("%-34s   last warp: issue %6.1f, complete %6.1f\n", "", h[(WARPS - 1) * 2] / 256.0, h[(WARPS - 1) * 2 + 1] / 256.0);
    cudaFree(d);
}

int main() {
    run<64, 1, 1, true>("dependent chain");
    run<64, 2, 1, true>("2 accumulators round-robin");
    run<64, 4, 1, true>("4 accumulators round-robin");
    run<16, 1, 1, true>("dependent chain");
    run<16, 4, 1, true>("4 accumulators round-robin");
    run<128, 1, 1, true>("dependent chain");
    run<256, 1, 1, true>("dependent chain");
    run<64, 1, 2, true>("2 warps, one chain each");
    run<64, 1, 4, true>("4 warps, one chain each");
    run<64, 2, 4, true>("4 warps, 2 accumulators each");
    run<64, 1, 1, false>("dependent chain");
    run<64, 4, 1, false>("4 accumulators round-robin");
    run<64, 1, 4, false>("4 warps, one chain each");
    return 0;
}
