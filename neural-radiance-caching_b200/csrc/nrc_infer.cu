// Fused NRC cache query for sm_100a: raw 60-byte RadianceQuery records in, float3 radiance out.
// Replaces the reference chain  CompositeEncoding::forward_impl -> kernel_mlp_fused<64,...,INFERENCE> ->
// trim_and_cast  (tcnn/src/fully_fused_mlp.cu:499-557, nrc/src/NRCNetwork.cu:64-83)
// with ONE persistent kernel: encoded features and activations never touch HBM.
//
// CTA = 4 MMA-issue warps + 4 groups x 4 worker warps. Each group owns one 128-query tile at a time:
//   workers : bulk-copy the tile's 7680 input bytes to smem, encode in registers, write the fp16 features
//             into TMEM as the A operand, then per layer read the fp32 accumulator from TMEM, ReLU, round to
//             fp16 and write it back to TMEM as the next layer's A operand.
//   MMA warp g: one elected thread issues tcgen05.mma.kind::f16 (M=128 queries, N=64|16, K=16) with A from TMEM
//             and the layer's weight matrix (resident in smem for the whole kernel) as B; completion is
//             signalled to the group through tcgen05.commit -> mbarrier.
// Four tiles are in flight per SM so the tensor pipe, the TMEM ports and the FP32 pipes overlap.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include "nrc_kernels.h"
#include "nrc_encode.cuh"

namespace nrcb {

constexpr int kGroups = 4;
constexpr int kTile = 128;
constexpr int kThreads = kGroups * 32 + kGroups * 128;  // 4 issuer warps + 16 worker warps
constexpr uint32_t kTmemGroupCols = 128;  // per group: D at +0 (64 cols), A at +64 (40 cols)
constexpr uint32_t kTmemA = 64;

template <int E>
struct InferSmem {
    static constexpr int kWin = 64 * E * 2;           // W_in image
    static constexpr int kWhid = 4 * 64 * 64 * 2;     // 4 hidden images
    static constexpr int kWout = 16 * 64 * 2;
    static constexpr int kWeights = kWin + kWhid + kWout;
    static constexpr int kStage = kTile * 15 * 4;     // 7680 B of raw queries per group
    static constexpr int kOut = kTile * 3 * 4;        // 1536 B of results per group
    static constexpr int kInOff = kWeights;
    static constexpr int kOutOff = kInOff + kGroups * kStage;
    static constexpr int kAcc = kTile * 12 + kTile * 16;   // fused accumulation: throughput (1536 B) + output (2048 B) of a tile
    static constexpr int kAccOff = kOutOff + kGroups * kOut;           // [kGroups][2 parities][kAcc]
    static constexpr int kBarOff = kAccOff + kGroups * 2 * kAcc;
    static constexpr int kTotal = kBarOff + 256;
};

// Copies one row-major fp16 matrix W[n_out][n_in] into the UMMA K-major interleaved image:
// off(o,k) = (k/8)*(n_out*16) + (o/8)*128 + (o%8)*16 + (k%8)*2     (8x8 core matrices, 128 B each)
__device__ __forceinline__ void stage_weight(uint8_t* dst, const __half* __restrict__ W, int n_out, int n_in, int tid,
                                             int nthr) {
    const int chunks = n_in / 8;
    for (int i = tid; i < n_out * chunks; i += nthr) {
        const int o = i / chunks, c = i - o * chunks;
        const uint4 v = *reinterpret_cast<const uint4*>(W + (size_t)o * n_in + c * 8);
        *reinterpret_cast<uint4*>(dst + c * (n_out * 16) + (o >> 3) * 128 + (o & 7) * 16) = v;
    }
}

__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, 128;" ::"r"(g + 1) : "memory"); }

// MMA issuer of group G. Everything the MMAs consume is warp-uniform and derived from compile-time group /
// layer indices, so ptxas keeps the operands in uniform registers and emits back-to-back UTCHMMA (a
// data-dependent scheduler here cost ~150 cycles per MMA in R2UR + elect waterfalls; measured, profiles/).
template <int ENC, int G>
__device__ __forceinline__ void mma_issuer(uint8_t* smem, uint32_t tmem_base, uint64_t* a_ready, uint64_t* d_ready,
                                           uint32_t n_tiles, uint32_t tile_stride) {
    constexpr int E = ENC == 0 ? 80 : 64;
    constexpr int KS0 = E / 16;
    using L = InferSmem<E>;
    constexpr uint32_t idesc64 = umma_idesc_f16(128, 64, 0, 0);
    constexpr uint32_t idesc16 = umma_idesc_f16(128, 16, 0, 0);
    const uint32_t sbase = smem_u32(smem);
    const uint64_t d_in = umma_desc(sbase, 64 * 16, 128);
    const uint64_t d_hid = umma_desc(sbase + L::kWin, 64 * 16, 128);
    const uint64_t d_out = umma_desc(sbase + L::kWin + L::kWhid, 16 * 16, 128);
    const uint32_t dD = tmem_base + G * kTmemGroupCols;
    const uint32_t dA = dD + kTmemA;
    const uint32_t first = blockIdx.x * kGroups + G;
    const uint32_t nt = first < n_tiles ? (n_tiles - first + tile_stride - 1) / tile_stride : 0;
    const bool leader = elect_one();
    uint32_t ph = 0;
    for (uint32_t it = 0; it < nt; ++it) {
#pragma unroll
        for (int ly = 0; ly < 6; ++ly) {
            mbar_wait_uniform(&a_ready[G], ph);
            ph ^= 1;
            tc_fence_after();
            if (leader) {
                if (ly == 0) {
#pragma unroll
                    for (int k = 0; k < KS0; ++k) umma_ts(dD, dA + 8 * k, d_in + (uint64_t)(k * 128), idesc64, k > 0);
                } else if (ly < 5) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_ts(dD, dA + 8 * k, d_hid + (uint64_t)((ly - 1) * 512 + k * 128), idesc64, k > 0);
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_ts(dD, dA + 8 * k, d_out + (uint64_t)(k * 32), idesc16, k > 0);
                }
                umma_commit(&d_ready[G]);
            }
            __syncwarp();
        }
    }
}

template <int ENC>  // 0 = Frequency (E = 80), 1 = Hash (E = 64)
__global__ void __launch_bounds__(kThreads, 1)
nrc_infer_kernel(const float* __restrict__ in, float* __restrict__ out, uint32_t n, const __half* __restrict__ params,
                 GridMeta gm, AccumArgs acc) {
    constexpr int E = ENC == 0 ? 80 : 64;
    using L = InferSmem<E>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    uint64_t* in_full = bars;              // [kGroups] raw tile landed in smem
    uint64_t* a_ready = bars + kGroups;    // [kGroups] A operand (and D drained) ready for the next layer's MMA
    uint64_t* d_ready = bars + 2 * kGroups;  // [kGroups] accumulator complete
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * kGroups);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t n_tiles = (n + kTile - 1) / kTile;
    const uint32_t tile_stride = gridDim.x * kGroups;

    // ---- one-time setup: weights -> smem images, barriers, TMEM
    {
        const __half* W = params;
        stage_weight(smem, W, 64, E, tid, kThreads);
        W += 64 * E;
        for (int l = 0; l < 4; ++l, W += 4096) stage_weight(smem + L::kWin + l * 8192, W, 64, 64, tid, kThreads);
        stage_weight(smem + L::kWin + L::kWhid, W, 16, 64, tid, kThreads);
    }
    if (tid == 0) {
        for (int g = 0; g < kGroups; ++g) { mbar_init(&in_full[g], 1); mbar_init(&a_ready[g], 4); mbar_init(&d_ready[g], 1); }
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < kGroups) {
        // ======================= MMA issuers: warp g serves group g =======================
        if (warp == 0) mma_issuer<ENC, 0>(smem, tmem_base, a_ready, d_ready, n_tiles, tile_stride);
        else if (warp == 1) mma_issuer<ENC, 1>(smem, tmem_base, a_ready, d_ready, n_tiles, tile_stride);
        else if (warp == 2) mma_issuer<ENC, 2>(smem, tmem_base, a_ready, d_ready, n_tiles, tile_stride);
        else mma_issuer<ENC, 3>(smem, tmem_base, a_ready, d_ready, n_tiles, tile_stride);
    } else {
        // ======================= workers =======================
        const int g = (warp - kGroups) >> 2;
        const int gtid = tid - kGroups * 32 - g * 128;            // 0..127 within the group
        const uint32_t row = (warp & 3) * 32 + lane;    // TMEM lane == query row inside the tile
        const uint32_t tD = tmem_base + g * kTmemGroupCols + (((warp & 3) * 32u) << 16);
        const uint32_t tA = tD + kTmemA;
        float* s_in = reinterpret_cast<float*>(smem + L::kInOff + g * L::kStage);
        float* s_out = reinterpret_cast<float*>(smem + L::kOutOff + g * L::kOut);
        const bool in_bulk_ok = (reinterpret_cast<uintptr_t>(in) & 15) == 0;
        const bool out_vec_ok = (reinterpret_cast<uintptr_t>(out) & 15) == 0;

        uint8_t* s_acc = smem + L::kAccOff + g * 2 * L::kAcc;
        const bool acc_bulk_ok = acc.out != nullptr && (reinterpret_cast<uintptr_t>(acc.throughput) & 15) == 0 &&
                                 (reinterpret_cast<uintptr_t>(acc.out) & 15) == 0;
        // tiles whose 128 queries are all pixels prefetch throughput + output rows with the same mbarrier as the input
        auto acc_staged = [&](uint32_t t) { return acc_bulk_ok && in_bulk_ok && (t + 1) * kTile <= acc.n && (t + 1) * kTile <= n; };

        auto load_tile = [&](uint32_t t, uint32_t par) {  // called by all 128 threads of the group, after a group_sync
            const uint32_t rows = min((uint32_t)kTile, n - t * kTile);
            const float* src = in + (size_t)t * kTile * 15;
            if (in_bulk_ok && rows == kTile) {
                if (gtid == 0) {
                    const bool st = acc_staged(t);
                    mbar_arrive_expect_tx(&in_full[g], L::kStage + (st ? L::kAcc : 0));
                    bulk_g2s(s_in, src, L::kStage, &in_full[g]);
                    if (st) {
                        bulk_g2s(s_acc + par * L::kAcc, acc.throughput + (size_t)t * kTile * 3, kTile * 12, &in_full[g]);
                        bulk_g2s(s_acc + par * L::kAcc + kTile * 12, acc.out + (size_t)t * kTile, kTile * 16, &in_full[g]);
                    }
                }
            } else {  // ragged tail / unaligned caller buffer: plain loads, zero-fill the missing rows
                for (uint32_t i = gtid; i < kTile * 15; i += 128) s_in[i] = i < rows * 15 ? __ldg(src + i) : 0.0f;
                group_sync(g);
                if (gtid == 0) mbar_arrive(&in_full[g]);
            }
        };

        uint32_t t = blockIdx.x * kGroups + g;
        uint32_t ph_in = 0, ph_d = 0, par = 0;
        if (t < n_tiles) load_tile(t, 0);
        for (; t < n_tiles; t += tile_stride) {
            mbar_wait(&in_full[g], ph_in); ph_in ^= 1;
            float q[15];
#pragma unroll
            for (int i = 0; i < 15; ++i) q[i] = s_in[row * 15 + i];
            fence_proxy_async_smem();                        // order these generic reads before the next bulk write
            group_sync(g);                                   // everyone holds its query: the stage is free
            if (t + tile_stride < n_tiles) load_tile(t + tile_stride, par ^ 1);

            // ---- encode -> A operand of the input layer
            {
                uint32_t e[E / 2];
                if constexpr (ENC == 0) encode_frequency(q, e);
                else encode_hash(q, reinterpret_cast<const __half2*>(params + (64 * E + 4 * 4096 + 16 * 64)), gm, e);
#pragma unroll
                for (int c = 0; c < E / 16; ++c) {
                    uint32_t v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = e[c * 8 + i];
                    tmem_st8(tA + c * 8, v);
                }
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&a_ready[g]);

            // ---- 5 hidden layers: D (fp32) -> relu -> fp16 -> A
#pragma unroll 1
            for (int l = 0; l < 5; ++l) {
                mbar_wait(&d_ready[g], ph_d); ph_d ^= 1;
                tc_fence_after();
#pragma unroll
                for (int hcol = 0; hcol < 2; ++hcol) {
                    uint32_t v[32], w[16];
                    tmem_ld32(tD + hcol * 32, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 16; ++i) w[i] = pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
                    tmem_st16(tA + hcol * 16, w);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&a_ready[g]);
            }
            // ---- output layer: relu, fp16 rounding (the reference's output is half), first 3 of 16 lanes
            mbar_wait(&d_ready[g], ph_d); ph_d ^= 1;
            tc_fence_after();
            float y[3];
            {
                uint32_t v[16];
                tmem_ld16(tD, v);
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < 3; ++c) y[c] = __half2float(__float2half_rn(fmaxf(__uint_as_float(v[c]), 0.0f)));
            }
            // D is drained; the next tile's first arrive (after its encode) re-arms the MMA warp.
            const uint32_t rows = min((uint32_t)kTile, n - t * kTile);
            // Fused accumulate_render_radiance (nrc_helpers.cu:77-129) for the pixel queries: the radiance is still in
            // registers, so the separate pass (re-read 12 B radiance, 12 B throughput, 16 B output per pixel) disappears.
            if (acc.out != nullptr) {
                const uint32_t px = t * kTile + row;
                if (px < acc.n && row < rows) {
                    float t0, t1, t2;
                    float4 d;
                    if (acc_staged(t)) {
                        const float* st = reinterpret_cast<const float*>(s_acc + par * L::kAcc);
                        t0 = st[3 * row]; t1 = st[3 * row + 1]; t2 = st[3 * row + 2];
                        d = reinterpret_cast<const float4*>(s_acc + par * L::kAcc + kTile * 12)[row];
                    } else {
                        t0 = __ldg(acc.throughput + 3 * (size_t)px); t1 = __ldg(acc.throughput + 3 * (size_t)px + 1);
                        t2 = __ldg(acc.throughput + 3 * (size_t)px + 2);
                        d = acc.mode == 0 ? acc.out[px] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                    }
                    float4* o = acc.out + px;
                    switch (acc.mode) {
                        case 0: *o = make_float4(fmaf(t0 * y[0], acc.weight, d.x), fmaf(t1 * y[1], acc.weight, d.y), fmaf(t2 * y[2], acc.weight, d.z), 1.0f); break;
                        case 2: *o = make_float4(y[0] * t0, y[1] * t1, y[2] * t2, 1.0f); break;
                        case 4: *o = make_float4(y[0], y[1], y[2], 1.0f); break;
                        case 5: *o = make_float4(t0, t1, t2, 1.0f); break;
                        default: break;
                    }
                }
            }
            par ^= 1;
            float* dst = out + (size_t)t * kTile * 3;
            if (out_vec_ok && rows == kTile) {
                s_out[row * 3 + 0] = y[0]; s_out[row * 3 + 1] = y[1]; s_out[row * 3 + 2] = y[2];
                group_sync(g);
                if (gtid < 96) reinterpret_cast<float4*>(dst)[gtid] = reinterpret_cast<const float4*>(s_out)[gtid];
            } else if (row < rows) {
                dst[row * 3 + 0] = y[0]; dst[row * 3 + 1] = y[1]; dst[row * 3 + 2] = y[2];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 512);
}

template <int ENC>
static cudaError_t launch_infer_t(const float* in, float* out, uint32_t n, const __half* params, const GridMeta& gm,
                                  int num_sms, cudaStream_t stream, const AccumArgs& acc) {
    using L = InferSmem<ENC == 0 ? 80 : 64>;
    static bool configured[2] = {false, false};
    if (!configured[ENC]) {
        cudaError_t e = cudaFuncSetAttribute(nrc_infer_kernel<ENC>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
        if (e != cudaSuccess) return e;
        configured[ENC] = true;
    }
    const uint32_t n_tiles = (n + kTile - 1) / kTile;
    const uint32_t grid = min((uint32_t)num_sms, (n_tiles + kGroups - 1) / kGroups);
    nrc_infer_kernel<ENC><<<grid, kThreads, L::kTotal, stream>>>(in, out, n, params, gm, acc);
    return cudaGetLastError();
}

cudaError_t launch_infer(int encoding, const float* in, float* out, uint32_t n, const __half* params, const GridMeta& gm,
                         int num_sms, cudaStream_t stream, const AccumArgs* acc) {
    if (n == 0) return cudaSuccess;
    const AccumArgs a = acc ? *acc : AccumArgs{nullptr, nullptr, 0, 0, 0.0f};
    return encoding == 0 ? launch_infer_t<0>(in, out, n, params, gm, num_sms, stream, a)
                         : launch_infer_t<1>(in, out, n, params, gm, num_sms, stream, a);
}

}  // namespace nrcb
