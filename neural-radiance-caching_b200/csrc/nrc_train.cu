// Fused NRC training step for sm_100a.
//   kernel 1 (nrc_train_tile_kernel): one CTA per 128-sample tile does what the reference spreads over ~20
//     launches: encodings, forward MLP with activations kept in shared memory,
//     RelativeL2Luminance loss + dL/dy (tcnn/losses/relative_l2_luminance.h:54-86), output-ReLU transfer,
//     the backward chain (tcnn/src/fully_fused_mlp.cu:150-259) and the per-tile weight gradients
//     dW = dY * X^T (fully_fused_mlp.cu:785,819,828) -- every contraction on tcgen05, accumulators in TMEM.
//   kernel 2 (nrc_reduce_grads_kernel): deterministic sum of the per-tile partials -> fp32 gradient vector.
//   kernel 3 (nrc_adam_ema_kernel): fp16 gradient, Adam (tcnn/optimizers/adam.h:75-118), EMA (ema.h:62-76).
//
// Shared-memory operands use the SWIZZLE_NONE interleaved layout (8x8 fp16 core matrices of 128 B). One image of a
// [rows x cols] matrix serves both as a K-major operand (K = cols) and, read through an MN-major descriptor,
// as the transposed operand (K = rows) -- so W is staged once for forward and backward, and each activation /
// delta tile is written once and consumed by the next layer's MMA and by the weight-gradient MMA.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include "nrc_kernels.h"
#include "nrc_encode.cuh"

namespace nrcb {

constexpr int kTT = 128;             // samples per tile
constexpr int kTrainThreads = 160;   // 4 worker warps (one sample per thread) + 1 MMA-issue warp

template <int E>
struct TrainSmem {
    static constexpr int kWin = 64 * E * 2, kWhid = 4 * 8192, kWout = 2048;
    static constexpr int kW = 0;                              // weight images: W_in | W_h1..4 | W_out
    static constexpr int kX = kW + kWin + kWhid + kWout;      // encoded inputs   [128 x E]
    static constexpr int kH = kX + 128 * E * 2;               // h1..h5           5 x [128 x 64]
    static constexpr int kDelta = kH + 5 * 16384;             // delta ping-pong  2 x [128 x 64]
    static constexpr int kG = kDelta + 2 * 16384;             // dL/dy            [128 x 16]
    static constexpr int kIn = kG + 4096;                     // raw queries 7680 B
    static constexpr int kBar = kIn + 7680;
    static constexpr int kTotal = kBar + 128;
};

// TMEM columns
constexpr uint32_t kAcc = 0;      // forward / backward accumulator (64 cols)
constexpr uint32_t kWgIn = 64;    // dW_in   [64 x E]   (M = 64: row r lives in lane (r/16)*32 + r%16)
__host__ __device__ constexpr uint32_t wg_hid_col(int E, int l) { return 64 + E + 64 * l; }  // dW_h(l+1), l = 0..3
__host__ __device__ constexpr uint32_t wg_out_col(int E) { return 64 + E + 256; }            // dW_out^T [64 x 16]

__device__ __forceinline__ void stage_weight_t(uint8_t* dst, const __half* __restrict__ W, int n_out, int n_in, int tid,
                                               int nthr) {
    const int chunks = n_in / 8;
    for (int i = tid; i < n_out * chunks; i += nthr) {
        const int o = i / chunks, c = i - o * chunks;
        const uint4 v = *reinterpret_cast<const uint4*>(W + (size_t)o * n_in + c * 8);
        *reinterpret_cast<uint4*>(dst + c * (n_out * 16) + (o >> 3) * 128 + (o & 7) * 16) = v;
    }
}

// row r of a [128 x cols] interleaved image: 16-byte chunk c lives at c*2048 + (r/8)*128 + (r%8)*16
__device__ __forceinline__ uint8_t* row_chunk(uint8_t* img, uint32_t r, uint32_t c) {
    return img + c * 2048 + (r >> 3) * 128 + (r & 7) * 16;
}

// ---- MMA issue sequences (one elected thread). Descriptor strides: see header comment / sm100_prims.cuh.
// K-major A [128 x K] (lbo 2048, sbo 128, +4096 per k-step); K-major B = weight image [N x K] (lbo N*16);
// MN-major views: activation images (sbo 2048, lbo 128, +256 per k-step of 16 samples),
//                 weight images as W^T (sbo = n_out*16, lbo 128, +256 per k-step of 16 outputs).
__device__ __forceinline__ void mma_fwd(uint32_t d, uint32_t a_img, uint32_t w_img, int n_out, int ksteps) {
    const uint32_t idesc = umma_idesc_f16(128, n_out, 0, 0);
#pragma unroll
    for (int k = 0; k < 5; ++k)
        if (k < ksteps) umma_ss(d, umma_desc(a_img + k * 4096, 2048, 128), umma_desc(w_img + k * 2 * n_out * 16, n_out * 16, 128), idesc, k > 0);
}
// D[128 x n] = A[128 x K=n_out] * W[n_out x n]      (W image read MN-major)
__device__ __forceinline__ void mma_dgrad(uint32_t d, uint32_t a_img, uint32_t w_img, int n_out, int n) {
    const uint32_t idesc = umma_idesc_f16(128, n, 0, 1);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (k < n_out / 16) umma_ss(d, umma_desc(a_img + k * 4096, 2048, 128), umma_desc(w_img + k * 256, 128, n_out * 16), idesc, k > 0);
}
// D[64 x n] = Dl^T[64 x 128] * Act[128 x n]         (both images read MN-major, K = 128 samples)
__device__ __forceinline__ void mma_wgrad(uint32_t d, uint32_t delta_img, uint32_t act_img, int n) {
    const uint32_t idesc = umma_idesc_f16(64, n, 1, 1);
#pragma unroll
    for (int k = 0; k < 8; ++k)
        umma_ss(d, umma_desc(delta_img + k * 256, 128, 2048), umma_desc(act_img + k * 256, 128, 2048), idesc, k > 0);
}

template <int ENC>
__global__ void __launch_bounds__(kTrainThreads, 1)
nrc_train_tile_kernel(const float* __restrict__ in, const float* __restrict__ targets, uint32_t n_global,
                      const __half* __restrict__ params, float* __restrict__ partials, float* __restrict__ grid_grad,
                      float* __restrict__ loss_partials, GridMeta gm) {
    constexpr int E = ENC == 0 ? 80 : 64;
    constexpr int NMLP = 64 * E + 4 * 4096 + 1024;
    using L = TrainSmem<E>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBar);
    uint64_t* a_ready = bars;       // workers -> issuer (count 4)
    uint64_t* d_ready = bars + 1;   // issuer  -> workers
    uint64_t* in_full = bars + 2;   // bulk copy of the raw queries
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
    float* warp_loss = reinterpret_cast<float*>(bars + 4);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t tile = blockIdx.x;

    {
        const __half* W = params;
        stage_weight_t(smem + L::kW, W, 64, E, tid, kTrainThreads);
        W += 64 * E;
        for (int l = 0; l < 4; ++l, W += 4096) stage_weight_t(smem + L::kW + L::kWin + l * 8192, W, 64, 64, tid, kTrainThreads);
        stage_weight_t(smem + L::kW + L::kWin + L::kWhid, W, 16, 64, tid, kTrainThreads);
    }
    if (tid == 0) {
        mbar_init(a_ready, 4); mbar_init(d_ready, 1); mbar_init(in_full, 1);
        fence_mbar_init();
        const float* src = in + (size_t)tile * kTT * 15;
        if ((reinterpret_cast<uintptr_t>(in) & 15) == 0) {
            mbar_arrive_expect_tx(in_full, 7680);
            bulk_g2s(smem + L::kIn, src, 7680, in_full);
        }
    }
    if (warp == 4) tmem_alloc(tmem_slot, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t sW = smem_u32(smem + L::kW), sX = smem_u32(smem + L::kX), sH = smem_u32(smem + L::kH);
    const uint32_t sDl = smem_u32(smem + L::kDelta), sG = smem_u32(smem + L::kG);

    if (warp == 4) {
        // ======================= MMA issuer =======================
        const bool leader = elect_one();
        uint32_t ph = 0;
        auto wait_workers = [&]() { mbar_wait_uniform(a_ready, ph); ph ^= 1; tc_fence_after(); };
        const uint32_t acc = tmem_base + kAcc;
        // forward
        wait_workers();
        if (leader) { mma_fwd(acc, sX, sW, 64, E / 16); umma_commit(d_ready); }
        __syncwarp();
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            wait_workers();
            if (leader) { mma_fwd(acc, sH + l * 16384, sW + L::kWin + l * 8192, 64, 4); umma_commit(d_ready); }
            __syncwarp();
        }
        wait_workers();
        if (leader) { mma_fwd(acc, sH + 4 * 16384, sW + L::kWin + L::kWhid, 16, 4); umma_commit(d_ready); }
        __syncwarp();
        // backward: dL/dy (G) ready
        wait_workers();
        if (leader) {
            mma_dgrad(acc, sG, sW + L::kWin + L::kWhid, 16, 64);            // (W_out^T g)
            umma_commit(d_ready);
            mma_wgrad(tmem_base + wg_out_col(E), sH + 4 * 16384, sG, 16);     // dW_out^T = h5^T g
        }
        __syncwarp();
#pragma unroll
        for (int l = 3; l >= 0; --l) {   // delta_(l+2) ready in buffer (3-l)&1 ; produces delta_(l+1)
            wait_workers();
            if (leader) {
                const uint32_t dl = sDl + ((3 - l) & 1) * 16384;
                mma_dgrad(acc, dl, sW + L::kWin + l * 8192, 64, 64);
                umma_commit(d_ready);
                mma_wgrad(tmem_base + wg_hid_col(E, l), dl, sH + l * 16384, 64);  // dW_h(l+1) = delta_(l+2)^T h_(l+1)
            }
            __syncwarp();
        }
        wait_workers();                   // delta_1 ready in buffer 0
        if (leader) {
            if (ENC == 1) mma_dgrad(acc, sDl, sW, 64, 32);                      // dL/d(grid features)
            mma_wgrad(tmem_base + kWgIn, sDl, sX, E);                           // dW_in = delta_1^T x
            umma_commit(d_ready);
        }
        __syncwarp();
    } else {
        // ======================= workers: thread = sample =======================
        const uint32_t row = warp * 32 + lane;
        const uint32_t tq = tmem_base + ((warp * 32u) << 16);
        const uint32_t n_total = n_global * 3;
        uint32_t ph = 0;
        auto wait_mma = [&]() { mbar_wait(d_ready, ph); ph ^= 1; tc_fence_after(); };
        auto signal = [&]() { fence_proxy_async_smem(); tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(a_ready); };

        float q[15];
        if ((reinterpret_cast<uintptr_t>(in) & 15) == 0) {
            mbar_wait(in_full, 0);
            const float* s_in = reinterpret_cast<const float*>(smem + L::kIn);
#pragma unroll
            for (int i = 0; i < 15; ++i) q[i] = s_in[row * 15 + i];
        } else {
#pragma unroll
            for (int i = 0; i < 15; ++i) q[i] = __ldg(in + ((size_t)tile * kTT + row) * 15 + i);
        }
        float tg[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) tg[c] = __ldg(targets + ((size_t)tile * kTT + row) * 3 + c);
        {
            uint32_t e[E / 2];
            if constexpr (ENC == 0) encode_frequency(q, e);
            else encode_hash(q, reinterpret_cast<const __half2*>(params + NMLP), gm, e);
#pragma unroll
            for (int c = 0; c < E / 8; ++c)
                *reinterpret_cast<uint4*>(row_chunk(smem + L::kX, row, c)) = make_uint4(e[4 * c], e[4 * c + 1], e[4 * c + 2], e[4 * c + 3]);
        }
        signal();
        // forward: h_(l+1) = relu(acc) -> fp16 -> smem; keep the relu mask (2 x 32 bits per layer) in registers
        uint32_t mask[5][2];
#pragma unroll
        for (int l = 0; l < 5; ++l) {
            wait_mma();
#pragma unroll
            for (int hc = 0; hc < 2; ++hc) {
                uint32_t v[32];
                tmem_ld32(tq + kAcc + hc * 32, v);
                tmem_wait_ld();
                uint32_t m = 0, w[16];
#pragma unroll
                for (int i = 0; i < 32; ++i) m |= (__uint_as_float(v[i]) > 0.0f ? 1u : 0u) << i;
#pragma unroll
                for (int i = 0; i < 16; ++i) w[i] = pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
                mask[l][hc] = m;
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    *reinterpret_cast<uint4*>(row_chunk(smem + L::kH + l * 16384, row, hc * 4 + c)) =
                        make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
            }
            signal();
        }
        // output + loss (relative_l2_luminance.h:62-86) + output-ReLU transfer (common_device.h:241-249)
        wait_mma();
        float loss = 0.0f;
        {
            uint32_t v[16];
            tmem_ld16(tq + kAcc, v);
            tmem_wait_ld();
            float y[3], gr[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) y[c] = __half2float(__float2half_rn(fmaxf(__uint_as_float(v[c]), 0.0f)));
            const float lum = 0.299f * y[0] + 0.587f * y[1] + 0.114f * y[2];
            const float den = lum * lum + 0.01f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float diff = y[c] - tg[c];
                loss += diff * diff / den / (float)n_total;
                const float gh = __half2float(__float2half_rn(128.0f * (2.0f * diff / den) / (float)n_total));
                gr[c] = y[c] > 0.0f ? gh : 0.0f;
            }
            *reinterpret_cast<uint4*>(row_chunk(smem + L::kG, row, 0)) = make_uint4(pack_h2(gr[0], gr[1]), pack_h2(gr[2], 0.0f), 0u, 0u);
            *reinterpret_cast<uint4*>(row_chunk(smem + L::kG, row, 1)) = make_uint4(0u, 0u, 0u, 0u);
        }
        signal();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
        if (lane == 0) warp_loss[warp] = loss;
        // backward chain: delta_l = (W^T delta_(l+1)) .* [h_l > 0], rounded to fp16
#pragma unroll
        for (int l = 4; l >= 0; --l) {
            wait_mma();
            uint8_t* dbuf = smem + L::kDelta + ((4 - l) & 1) * 16384;
#pragma unroll
            for (int hc = 0; hc < 2; ++hc) {
                uint32_t v[32], w[16];
                tmem_ld32(tq + kAcc + hc * 32, v);
                tmem_wait_ld();
                const uint32_t m = mask[l][hc];
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    w[i] = pack_h2((m >> (2 * i)) & 1 ? __uint_as_float(v[2 * i]) : 0.0f,
                                   (m >> (2 * i + 1)) & 1 ? __uint_as_float(v[2 * i + 1]) : 0.0f);
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    *reinterpret_cast<uint4*>(row_chunk(dbuf, row, hc * 4 + c)) = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
            }
            signal();
        }
        // all MMAs done: weight-gradient accumulators (and, Hash, dL/d(grid features)) are in TMEM
        wait_mma();
        if (warp == 0 && lane == 0) loss_partials[tile] = (warp_loss[0] + warp_loss[1]) + (warp_loss[2] + warp_loss[3]);
        if constexpr (ENC == 1) {
            uint32_t v[32];
            tmem_ld32(tq + kAcc, v);
            tmem_wait_ld();
#pragma unroll 1
            for (int lv = 0; lv < 16; ++lv) {
                const float g0 = __half2float(__float2half_rn(__uint_as_float(v[2 * lv])));
                const float g1 = __half2float(__float2half_rn(__uint_as_float(v[2 * lv + 1])));
                if (g0 == 0.0f && g1 == 0.0f) continue;
                const uint32_t size = gm.offset[lv + 1] - gm.offset[lv], res = gm.resolution[lv];
                const bool dense = (gm.dense_mask >> lv) & 1;
                float p[3];
                uint32_t gp[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    p[d] = fmaf(gm.scale[lv], q[d], 0.5f);
                    const float fl = floorf(p[d]);
                    gp[d] = (uint32_t)(int)fl;
                    p[d] -= fl;
                }
                float* gl = grid_grad + (size_t)gm.offset[lv] * 2;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    float wgt = 1.0f;
                    uint32_t cc[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        if (c & (1 << d)) { wgt *= p[d]; cc[d] = gp[d] + 1; } else { wgt *= 1.0f - p[d]; cc[d] = gp[d]; }
                    }
                    const uint32_t idx = (dense ? cc[0] + cc[1] * res + cc[2] * res * res
                                                : cc[0] ^ (cc[1] * 2654435761u) ^ (cc[2] * 805459861u)) & (size - 1);
                    const float wh = __half2float(__float2half_rn(wgt));   // (GRAD_T)weight, grid.h:254
                    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * idx), "f"(wh * g0), "f"(wh * g1) : "memory");
                }
            }
        }
        // per-tile weight-gradient partials: M = 64 accumulators keep row o in lane (o/16)*32 + o%16.
        // The TMEM loads are warp-wide (.sync.aligned): every lane executes them, lanes 0..15 store.
        {
            const bool owner = lane < 16;
            const uint32_t o = warp * 16 + (lane & 15);
            float* P = partials + (size_t)tile * NMLP;
            auto dump = [&](uint32_t col, float* dst, int ncols) {   // dst: row o of a row-major [64 x ncols] block
                for (int c0 = 0; c0 < ncols; c0 += 16) {
                    uint32_t v[16];
                    tmem_ld16(tq + col + c0, v);
                    tmem_wait_ld();
                    if (owner) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            reinterpret_cast<float4*>(dst + c0)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                                 __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                    }
                }
            };
            dump(kWgIn, P + o * E, E);
#pragma unroll
            for (int l = 0; l < 4; ++l) dump(wg_hid_col(E, l), P + 64 * E + l * 4096 + o * 64, 64);
            {   // dW_out^T [64 x 16]: element (k = o, c) -> dW_out[c][k]
                uint32_t v[16];
                tmem_ld16(tq + wg_out_col(E), v);
                tmem_wait_ld();
                if (owner) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) P[64 * E + 4 * 4096 + c * 64 + o] = __uint_as_float(v[c]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- gradient reduction
// grad[i] = sum over tiles of partials[t][i] (fixed order -> bit-reproducible), grid part copied from the
// fp32 scatter accumulator; block 0 also reduces the per-tile losses (Trainer::loss, trainer.h:205-207).
__global__ void __launch_bounds__(256)
nrc_reduce_grads_kernel(const float* __restrict__ partials, const float* __restrict__ grid_grad,
                        const float* __restrict__ loss_partials, uint32_t n_tiles, uint32_t n_params, uint32_t n_mlp,
                        float* __restrict__ grad, float* __restrict__ loss_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_mlp) {
        float a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        uint32_t t = 0;
        for (; t + 4 <= n_tiles; t += 4) {
            a0 += partials[(size_t)(t + 0) * n_mlp + i];
            a1 += partials[(size_t)(t + 1) * n_mlp + i];
            a2 += partials[(size_t)(t + 2) * n_mlp + i];
            a3 += partials[(size_t)(t + 3) * n_mlp + i];
        }
        for (; t < n_tiles; ++t) a0 += partials[(size_t)t * n_mlp + i];
        grad[i] = (a0 + a1) + (a2 + a3);
    } else if (i < n_params) {
        grad[i] = grid_grad[i - n_mlp];
    }
    if (blockIdx.x == 0 && threadIdx.x < 32) {
        float l = 0;
        for (uint32_t t = threadIdx.x; t < n_tiles; t += 32) l += loss_partials[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
        if (threadIdx.x == 0) *loss_out = l;
    }
}

// ---------------------------------------------------------------- optimizer
// adam_step (tcnn/optimizers/adam.h:75-118) followed by ema_step_half_precision (ema.h:62-76) per parameter.
__global__ void __launch_bounds__(256)
nrc_adam_ema_kernel(OptimState os, const float* __restrict__ grad_f32, float* __restrict__ grid_grad_to_clear,
                    float ema_debias_old, float ema_debias_new) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= os.n_params) return;
    const __half gh = __float2half_rn(grad_f32[i]);   // the reference's gradient buffer is fp16 (trainer.h:322-336)
    os.grads[i] = gh;
    float gradient = __half2float(gh) / os.loss_scale;
    const bool is_matrix = i < os.n_mlp;
    if (!is_matrix) grid_grad_to_clear[i - os.n_mlp] = 0.0f;
    __half w_half = os.params[i];
    if (is_matrix || gradient != 0.0f) {
        const float weight_fp = os.master[i];
        if (is_matrix) gradient += os.l2_reg * weight_fp;
        const float gradient_sq = gradient * gradient;
        const float m1 = os.m1[i] = os.beta1 * os.m1[i] + (1.0f - os.beta1) * gradient;
        const float m2 = os.m2[i] = os.beta2 * os.m2[i] + (1.0f - os.beta2) * gradient_sq;
        const uint32_t cur = ++os.steps[i];
        const float lr = os.lr * (sqrtf(1.0f - powf(os.beta2, (float)cur)) / (1.0f - powf(os.beta1, (float)cur)));
        const float eff = fminf(fmaxf(lr / (sqrtf(m2) + os.eps), 0.0f), 3.402823466e+38f);
        const float new_weight = weight_fp - eff * m1;
        os.master[i] = new_weight;
        w_half = __float2half_rn(new_weight);
        os.params[i] = w_half;
    }
    const float filtered = (__half2float(os.ema[i]) * os.ema_decay * ema_debias_old +
                            __half2float(w_half) * (1.0f - os.ema_decay)) * ema_debias_new;
    os.ema[i] = __float2half_rn(filtered);
}

// ---------------------------------------------------------------- launchers
template <int ENC>
static cudaError_t launch_train_t(const float* in, const float* targets, uint32_t n_local, uint32_t n_global,
                                  const TrainBuffers& tb, const GridMeta& gm, cudaStream_t stream) {
    using L = TrainSmem<ENC == 0 ? 80 : 64>;
    static bool configured[2] = {false, false};
    if (!configured[ENC]) {
        cudaError_t e = cudaFuncSetAttribute(nrc_train_tile_kernel<ENC>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
        if (e != cudaSuccess) return e;
        configured[ENC] = true;
    }
    nrc_train_tile_kernel<ENC><<<n_local / kTT, kTrainThreads, L::kTotal, stream>>>(in, targets, n_global, tb.params, tb.partials,
                                                                                  tb.grid_grad, tb.loss_partials, gm);
    return cudaGetLastError();
}
cudaError_t launch_train_fwd_bwd(int encoding, const float* in, const float* targets, uint32_t n_local, uint32_t n_global,
                                 const TrainBuffers& tb, const GridMeta& gm, cudaStream_t stream) {
    return encoding == 0 ? launch_train_t<0>(in, targets, n_local, n_global, tb, gm, stream)
                         : launch_train_t<1>(in, targets, n_local, n_global, tb, gm, stream);
}
cudaError_t launch_reduce_grads(const TrainBuffers& tb, uint32_t n_tiles, uint32_t n_params, uint32_t n_mlp, float* grad_f32,
                                float* loss_out, cudaStream_t stream) {
    nrc_reduce_grads_kernel<<<(n_params + 255) / 256, 256, 0, stream>>>(tb.partials, tb.grid_grad, tb.loss_partials, n_tiles,
                                                                        n_params, n_mlp, grad_f32, loss_out);
    return cudaGetLastError();
}
cudaError_t launch_adam_ema(const OptimState& os, const float* grad_f32, float* grid_grad_to_clear, uint32_t step,
                            cudaStream_t stream) {
    // EmaOptimizer::step (ema.h:107-108): std::pow(float, unsigned) promotes to double
    const float debias_old = 1.0f - (float)pow((double)os.ema_decay, (double)(step - 1));
    const float debias_new = 1.0f / (1.0f - (float)pow((double)os.ema_decay, (double)step));
    nrc_adam_ema_kernel<<<(os.n_params + 255) / 256, 256, 0, stream>>>(os, grad_f32, grid_grad_to_clear, debias_old, debias_new);
    return cudaGetLastError();
}

}  // namespace nrcb
