// Fused NRC training step for sm_100a: ONE cooperative kernel per optimizer step.
//   phase 1 (per 128-sample tile, one CTA each): what the reference spreads over ~20 launches (SURVEY.md K1-K4,
//     K7-K13): encodings, forward MLP with activations kept in shared memory, RelativeL2Luminance loss + dL/dy
//     (tcnn/losses/relative_l2_luminance.h:54-86), output-ReLU transfer, the backward chain
//     (tcnn/src/fully_fused_mlp.cu:150-259) and the per-tile weight gradients dW = dY * X^T
//     (fully_fused_mlp.cu:785,819,828) -- every contraction on tcgen05, accumulators in TMEM.
//   phase 2: grid barrier (all CTAs are co-resident: cooperative launch, <= 128 CTAs of 1 per SM).
//   phase 3: CTA c sums slice c of the per-tile partials in fixed tile order (bit-reproducible) and applies
//     Adam (tcnn/optimizers/adam.h:75-118) + EMA (ema.h:62-76) to it -- replacing the reference's 6 split-K GEMM
//     reductions, adam_step, ema_step and the loss reduce_sum (K11, K14-K16).  With apply = 0 the reduced fp32
//     gradient is left in grad_f32 for a data-parallel all-reduce and nrc_adam_ema_kernel finishes the step.
//
// Shared-memory operands use the SWIZZLE_NONE interleaved layout (8x8 fp16 core matrices of 128 B). One image of a
// [rows x cols] matrix serves both as a K-major operand (K = cols) and, read through an MN-major descriptor,
// as the transposed operand (K = rows) -- so W is staged once for forward and backward, and each activation /
// delta tile is written once and consumed by the next layer's MMA and by the weight-gradient MMA.
// The weight image itself is kept up to date in global memory by the optimizer phase (params_img), so staging is a
// single bulk copy that overlaps the encode.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <type_traits>
#include "nrc_kernels.h"
#include "nrc_encode.cuh"
#include "nrc_frame.cuh"

#ifndef NRCB_TIMING_SKIP_WGRAD
#define NRCB_TIMING_SKIP_WGRAD 0
#endif
#ifndef NRCB_TIMING_SKIP_DUMP
#define NRCB_TIMING_SKIP_DUMP 0
#endif
#ifndef NRCB_TIMING_ADAM_FEWER_WRITES
#define NRCB_TIMING_ADAM_FEWER_WRITES 0
#endif
#ifndef NRCB_TIMING_ONLY_LEVEL
#define NRCB_TIMING_ONLY_LEVEL -1
#endif
#ifndef NRCB_TIMING_SKIP_FINE
#define NRCB_TIMING_SKIP_FINE 0
#endif
#ifndef NRCB_TIMING_SKIP_COARSE
#define NRCB_TIMING_SKIP_COARSE 0
#endif
namespace nrcb {

constexpr int kTT = 128;             // samples per tile
constexpr int kWorkerWarps = 8;      // warp w: TMEM lane quadrant w & 3 (rows), column half w >> 2
constexpr int kTrainThreads = (kWorkerWarps + 1) * 32;   // + 1 MMA-issue warp

template <int E>
struct TrainSmem {
    static constexpr int kWin = 64 * E * 2, kWhid = 4 * 8192, kWout = 2048;
    static constexpr int kWbytes = kWin + kWhid + kWout;
    static constexpr int kW = 0;                              // weight images: W_in | W_h1..4 | W_out
    static constexpr int kX = kW + kWbytes;                   // encoded inputs   [128 x E]
    static constexpr int kH = kX + 128 * E * 2;               // h1..h5           5 x [128 x 64]
    static constexpr int kDelta = kH + 5 * 16384;             // delta ping-pong  2 x [128 x 64]
    static constexpr int kG = kDelta + 2 * 16384;             // dL/dy            [128 x 16]
    static constexpr int kIn = kG + 4096;                     // raw queries 7680 B
    static constexpr int kBar = kIn + 7680;
    static constexpr int kTotal = kBar + 128;
    static constexpr int kScratch = kX;                       // phase 3 reuses the activation area: reduction scratch, lr table at + 8192
    // Hash optimizer phase: the CTA's slice of the grid optimizer state, bulk-copied in while the grid barrier is pending.
    static constexpr int kStage16 = kW;                       // [params | ema]            2 x n x 2 B  (the weight image is dead by then)
    static constexpr int kStage32 = kX + 8192 + 256;          // [master | m1 | m2 | steps] 4 x n x 4 B  (behind scratch + lr table)
    static constexpr int kStageMax = ((kBar - kStage32) / 16 < kWbytes / 4 ? (kBar - kStage32) / 16 : kWbytes / 4) & ~7;   // parameters
};

// TMEM columns
constexpr uint32_t kAcc = 0;      // forward / backward accumulator (64 cols)
constexpr uint32_t kWgIn = 64;    // dW_in   [64 x E]   (M = 64: row r lives in lane (r/16)*32 + r%16)
__host__ __device__ constexpr uint32_t wg_hid_col(int E, int l) { return 64 + E + 64 * l; }  // dW_h(l+1), l = 0..3
__host__ __device__ constexpr uint32_t wg_out_col(int E) { return 64 + E + 256; }            // dW_out^T [64 x 16]

// Position (in halves) of parameter i inside the UMMA weight image: matrix [n_out x n_in] row-major ->
// (k/8)*(n_out*8) + (o/8)*64 + (o%8)*8 + k%8   (8x8 core matrices of 64 halves).
__host__ __device__ __forceinline__ uint32_t img_index(uint32_t i, uint32_t E) {
    uint32_t base, n_out, n_in, r;
    if (i < 64 * E) { base = 0; n_out = 64; n_in = E; r = i; }
    else if (i < 64 * E + 4 * 4096) { r = i - 64 * E; base = 64 * E + (r >> 12) * 4096; r &= 4095; n_out = 64; n_in = 64; }
    else { base = 64 * E + 4 * 4096; r = i - base; n_out = 16; n_in = 64; }
    const uint32_t o = r / n_in, k = r - o * n_in;
    return base + (k >> 3) * (n_out * 8) + (o >> 3) * 64 + (o & 7) * 8 + (k & 7);
}

// row r of a [128 x cols] interleaved image: 16-byte chunk c lives at c*2048 + (r/8)*128 + (r%8)*16
__device__ __forceinline__ uint8_t* row_chunk(uint8_t* img, uint32_t r, uint32_t c) {
    return img + c * 2048 + (r >> 3) * 128 + (r & 7) * 16;
}

// ---- MMA issue sequences (one elected thread). Descriptor strides: see header comment / sm100_prims.cuh.
// K-major A [128 x K] (lbo 2048, sbo 128, +4096 per k-step); K-major B = weight image [N x K] (lbo N*16);
// MN-major views: activation images (sbo 2048, lbo 128, +256 per k-step of 16 samples),
//                 weight images as W^T (sbo = n_out*16, lbo 128, +256 per k-step of 16 outputs).
// acc16: fp16 accumulators (one fp16 per 32-bit TMEM column, rounded after every K = 16 instruction) -- the arithmetic of the
// reference's wmma::accumulator<__half> fragments in its forward and backward kernels (fully_fused_mlp.cu:68,150-259).
__device__ __forceinline__ void mma_fwd(uint32_t d, uint32_t a_img, uint32_t w_img, int n_out, int ksteps, bool acc16) {
    const uint32_t idesc = acc16 ? umma_idesc_f16_acc16(128, n_out, 0, 0) : umma_idesc_f16(128, n_out, 0, 0);
#pragma unroll
    for (int k = 0; k < 5; ++k)
        if (k < ksteps) umma_ss(d, umma_desc(a_img + k * 4096, 2048, 128), umma_desc(w_img + k * 2 * n_out * 16, n_out * 16, 128), idesc, k > 0);
}
// D[128 x n] = A[128 x K=n_out] * W[n_out x n]      (W image read MN-major)
__device__ __forceinline__ void mma_dgrad(uint32_t d, uint32_t a_img, uint32_t w_img, int n_out, int n, bool acc16) {
    const uint32_t idesc = acc16 ? umma_idesc_f16_acc16(128, n, 0, 1) : umma_idesc_f16(128, n, 0, 1);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (k < n_out / 16) umma_ss(d, umma_desc(a_img + k * 4096, 2048, 128), umma_desc(w_img + k * 256, 128, n_out * 16), idesc, k > 0);
}
// D[64 x n] = Dl^T[64 x 128] * Act[128 x n]         (both images read MN-major, K = 128 samples)
__device__ __forceinline__ void mma_wgrad(uint32_t d, uint32_t delta_img, uint32_t act_img, int n) {
#if NRCB_TIMING_SKIP_WGRAD      // timing experiment only (wrong gradients): what the weight-gradient MMAs cost the backward chain
    return;
#endif
    const uint32_t idesc = umma_idesc_f16(64, n, 1, 1);
#pragma unroll
    for (int k = 0; k < 8; ++k)
        umma_ss(d, umma_desc(delta_img + k * 256, 128, 2048), umma_desc(act_img + k * 256, 128, 2048), idesc, k > 0);
}

// The optimizer state of ONE matrix parameter, loaded before the grid barrier (it does not depend on the gradient), and
// the bias-corrected learning rate of its next step count (two powf) evaluated while the barrier is pending.
struct ParamState {
    float master, m1, m2, lr;
    uint32_t steps;
    __half param, ema;
};
__device__ __forceinline__ ParamState load_param_state(const TrainArgs& a, uint32_t i) {
    ParamState p;
    p.master = a.master[i]; p.m1 = a.m1[i]; p.m2 = a.m2[i]; p.steps = a.steps[i]; p.param = a.params[i]; p.ema = a.ema[i];
    const uint32_t cur = p.steps + 1;
    p.lr = a.lr * (sqrtf(1.0f - powf(a.beta2, (float)cur)) / (1.0f - powf(a.beta1, (float)cur)));
    return p;
}
// adam_step + ema_step_half_precision for a MATRIX parameter (i < n_mlp) on prefetched state.  Every product / sum is
// written with explicit round-to-nearest intrinsics so that this path and adam_ema_param (data-parallel path, stand-alone
// optimizer kernel) round identically whatever the compiler would contract: the replicas of a data-parallel job and a
// single-GPU run stay bit-identical.
__device__ __forceinline__ void adam_ema_matrix_param(const TrainArgs& a, uint32_t i, float g32, uint32_t E, uint32_t s, const ParamState& p) {
    const __half gh = __float2half_rn(g32);
    a.grads[i] = gh;
    const float gradient = fmaf(a.l2_reg, p.master, __half2float(gh) / a.loss_scale);
    const float m1 = fmaf(a.beta1, p.m1, __fmul_rn(1.0f - a.beta1, gradient));
    const float m2 = fmaf(a.beta2, p.m2, __fmul_rn(1.0f - a.beta2, __fmul_rn(gradient, gradient)));
    a.m1[i] = m1;
    a.m2[i] = m2;
    a.steps[i] = p.steps + 1;
    const float eff = fminf(fmaxf(p.lr / (sqrtf(m2) + a.eps), 0.0f), 3.402823466e+38f);
    const float new_weight = fmaf(-eff, m1, p.master);
    a.master[i] = new_weight;
    const __half w_half = __float2half_rn(new_weight);
    a.params[i] = w_half;
    const uint32_t img = img_index(i, E);
    a.params_img[img] = w_half;
    const float old_part = __fmul_rn(__fmul_rn(__half2float(p.ema), a.ema_decay), a.ema_debias_old[s]);
    const float filtered = __fmul_rn(fmaf(__half2float(w_half), 1.0f - a.ema_decay, old_part), a.ema_debias_new[s]);
    const __half e_half = __float2half_rn(filtered);
    a.ema[i] = e_half;
    a.ema_img[img] = e_half;        // the query kernel bulk-copies this image instead of re-formatting the weights per CTA
}

// ---------------------------------------------------------------- optimizer: ONE grid parameter, in registers
// adam_step for a parameter without weight decay (adam.h:75-118: skipped entirely when its gradient is zero) followed by
// ema_step_half_precision (ema.h:62-76).  Branch-free on purpose: the update is always evaluated and SELECTED by `nz`, so the
// compiler interleaves the independent chains of the parameters a thread holds (with a branch per parameter the IEEE division /
// square-root sequences ran one after the other: ~165 instructions per parameter on 9 warps per SM were the 12 us of this phase).
// Every product and sum is an explicit round-to-nearest intrinsic: the fused step kernel (both forms of its grid pass) and the
// stand-alone optimizer kernel of the NCCL path round identically whatever the compiler would contract.
// lr_tab (may be null): lr(common - d) for d < kLrTab, see grid_adam4.
constexpr uint32_t kLrTab = 64;   // entries of the per-step learning-rate table
__device__ __forceinline__ float adam_bias_corrected_lr(const TrainArgs& a, uint32_t cur) {
    return a.lr * (sqrtf(1.0f - powf(a.beta2, (float)cur)) / (1.0f - powf(a.beta1, (float)cur)));
}
__device__ __forceinline__ bool grid_param_update(const TrainArgs& a, float g32, uint32_t common, uint32_t s, const float* lr_tab,
                                                  float& w, float& m1, float& m2, uint32_t& st, __half& p, __half& e, __half& gh) {
    gh = __float2half_rn(g32);            // the reference's gradient buffer is fp16 (trainer.h:322-336)
    const float gf = __half2float(gh);
    const float gradient = a.loss_scale_pow2 ? __fmul_rn(gf, a.inv_loss_scale) : __fdiv_rn(gf, a.loss_scale);
    const bool nz = gradient != 0.0f;
    const float n1 = fmaf(a.beta1, m1, __fmul_rn(1.0f - a.beta1, gradient));
    const float n2 = fmaf(a.beta2, m2, __fmul_rn(1.0f - a.beta2, __fmul_rn(gradient, gradient)));
    const uint32_t cur = st + 1, missed = common - cur;
    float lr;
    if (lr_tab != nullptr && (missed < kLrTab || !nz)) lr = lr_tab[min(missed, kLrTab - 1)];   // (unused when !nz)
    else lr = adam_bias_corrected_lr(a, cur);
    const float eff = fminf(fmaxf(__fdiv_rn(lr, __fadd_rn(__fsqrt_rn(n2), a.eps)), 0.0f), 3.402823466e+38f);
    const float nw = fmaf(-eff, n1, w);
    m1 = nz ? n1 : m1;
    m2 = nz ? n2 : m2;
    st = nz ? cur : st;
    w = nz ? nw : w;
    p = nz ? __float2half_rn(nw) : p;
    const float old_part = __fmul_rn(__fmul_rn(__half2float(e), a.ema_decay), a.ema_debias_old[s]);
    e = __float2half_rn(__fmul_rn(fmaf(__half2float(p), 1.0f - a.ema_decay, old_part), a.ema_debias_new[s]));
    return nz;
}

// ---------------------------------------------------------------- optimizer (per parameter)
// adam_step (tcnn/optimizers/adam.h:75-118) followed by ema_step_half_precision (ema.h:62-76).  `g32` is the
// reduced fp32 gradient (loss-scaled); the reference's gradient buffer is fp16 (trainer.h:322-336), so it is
// rounded to fp16 first.  Also maintains the UMMA image of the fp16 training weights.
__device__ __forceinline__ void adam_ema_param(const TrainArgs& a, uint32_t i, float g32, uint32_t E, uint32_t s = 0) {
    if (i < a.n_mlp) { adam_ema_matrix_param(a, i, g32, E, s, load_param_state(a, i)); return; }
    float w = a.master[i], m1 = a.m1[i], m2 = a.m2[i];
    uint32_t st = a.steps[i];
    __half p = a.params[i], e = a.ema[i], gh;
    const bool nz = grid_param_update(a, g32, a.opt_step + s, s, nullptr, w, m1, m2, st, p, e, gh);
    a.grads[i] = gh;
    a.ema[i] = e;
    if (nz) { a.master[i] = w; a.m1[i] = m1; a.m2[i] = m2; a.steps[i] = st; a.params[i] = p; }
}

#ifndef NRCB_ADAM_STAGED
#define NRCB_ADAM_STAGED 0     // 1: Hash optimizer state of the CTA's slice staged in shared memory during the grid-barrier wait -- measured
                               // SLOWER (12.5 -> 13.2 us: the pass is bound by its arithmetic, not by its loads) and it makes the four-step kernel spill
#endif
#ifndef NRCB_ADAM_GROUPS
#define NRCB_ADAM_GROUPS 2      // 4 measured slower (15.4 vs 13.7 us, longer tail over the CTAs)
#endif
// Adam + EMA over grid parameters [glo, ghi) (multiples of 4) of one CTA, four parameters per thread and NRCB_ADAM_GROUPS groups
// in flight: the per-parameter version walked a dependent chain of ~7 L2 round trips per parameter (46 us measured).
// load_g(j) returns the reduced fp32 gradient of grid parameters j..j+3; clear: zero the scatter accumulator behind it.
// Adam + EMA for four consecutive grid parameters i .. i+3 (= grid parameters j .. j+3) whose gradient and state are in registers;
// writes the new state.  Shared by the two forms of the optimizer pass below, so that they round identically.
// The bias-corrected learning rate depends on the parameter's own step count `cur` only (adam.h:106), which is the step number
// `common` of this launch minus the steps whose batch missed the entry.  lr_tab[d] (shared memory, filled once per step by the CTA
// with the same expression) holds lr(common - d) for d < kLrTab: the two powf per parameter are left to entries that were missed
// more often than that.  (A batch misses ~2 % of the hashed entries, so after a few steps some lane of practically every warp
// needed the powf pair: ~4 us of the 16 us this phase took.)
struct GridState4 { float4 w, m1, m2; uint4 st; uint2 em, pr; };
template <class ClearG>
__device__ __forceinline__ void grid_adam4(const TrainArgs& a, uint32_t i, uint32_t j, const float4& g, const GridState4& v, uint32_t common, uint32_t s,
                                           const float* lr_tab, ClearG clear_g) {
    float gv[4] = {g.x, g.y, g.z, g.w}, wv[4] = {v.w.x, v.w.y, v.w.z, v.w.w};
    float av[4] = {v.m1.x, v.m1.y, v.m1.z, v.m1.w}, bv[4] = {v.m2.x, v.m2.y, v.m2.z, v.m2.w};
    uint32_t sv[4] = {v.st.x, v.st.y, v.st.z, v.st.w};
    __half ev[4], pv[4], gh[4];
    *reinterpret_cast<uint2*>(ev) = v.em;
    *reinterpret_cast<uint2*>(pv) = v.pr;
    bool any = false;
#pragma unroll
    for (int c = 0; c < 4; ++c) any |= grid_param_update(a, gv[c], common, s, lr_tab, wv[c], av[c], bv[c], sv[c], pv[c], ev[c], gh[c]);
#if !NRCB_TIMING_ADAM_FEWER_WRITES   // timing experiment only: the optimizer pass without its two gradient copies (6 of 30 B written per parameter)
    *reinterpret_cast<float4*>(a.grad_f32 + i) = g;
    *reinterpret_cast<uint2*>(a.grads + i) = *reinterpret_cast<uint2*>(gh);
#endif
    *reinterpret_cast<uint2*>(a.ema + i) = *reinterpret_cast<uint2*>(ev);
    if (any) {
        *reinterpret_cast<float4*>(a.master + i) = make_float4(wv[0], wv[1], wv[2], wv[3]);
        *reinterpret_cast<float4*>(a.m1 + i) = make_float4(av[0], av[1], av[2], av[3]);
        *reinterpret_cast<float4*>(a.m2 + i) = make_float4(bv[0], bv[1], bv[2], bv[3]);
        *reinterpret_cast<uint4*>(a.steps + i) = make_uint4(sv[0], sv[1], sv[2], sv[3]);
        *reinterpret_cast<uint2*>(a.params + i) = *reinterpret_cast<uint2*>(pv);
    }
    if (gv[0] != 0.0f || gv[1] != 0.0f || gv[2] != 0.0f || gv[3] != 0.0f) clear_g(j);   // also gradients that round to 0 in fp16
}
// Form 1 (any slice size): state and gradient from global memory, NRCB_ADAM_GROUPS groups in flight per thread.
template <class LoadG, class ClearG>
__device__ __forceinline__ void grid_adam_chunk(const TrainArgs& a, uint32_t base, uint32_t glo, uint32_t ghi, int tid, uint32_t E, uint32_t s, const float* lr_tab,
                                                LoadG load_g, ClearG clear_g) {
    const uint32_t common = a.opt_step + s;
    constexpr int kG = NRCB_ADAM_GROUPS;   // float4 groups of one thread in flight: 7 vector loads each (one L2 round trip for all)
    for (uint32_t j0 = glo + 4 * tid; j0 < ghi; j0 += 4 * kG * kTrainThreads) {
        float4 g[kG];
        GridState4 v[kG];
        bool live[kG];
#pragma unroll
        for (int u = 0; u < kG; ++u) {
            const uint32_t j = j0 + u * 4 * kTrainThreads;
            live[u] = j < ghi;
            if (live[u]) {
                const uint32_t i = base + j;
                g[u] = load_g(j);
                v[u].w = *reinterpret_cast<const float4*>(a.master + i);
                v[u].m1 = *reinterpret_cast<const float4*>(a.m1 + i);
                v[u].m2 = *reinterpret_cast<const float4*>(a.m2 + i);
                v[u].st = *reinterpret_cast<const uint4*>(a.steps + i);
                v[u].em = *reinterpret_cast<const uint2*>(a.ema + i);
                v[u].pr = *reinterpret_cast<const uint2*>(a.params + i);
            }
        }
#pragma unroll
        for (int u = 0; u < kG; ++u) {
            if (!live[u]) continue;
            const uint32_t j = j0 + u * 4 * kTrainThreads;
            grid_adam4(a, base + j, j, g[u], v[u], common, s, lr_tab, clear_g);
        }
    }
}
// Form 2: the CTA's whole state slice (20 B per parameter) was bulk-copied into shared memory while the grid barrier was pending
// (it does not depend on the gradient); after the barrier only the gradients come from L2 -- 4 B per parameter, the loads of
// kStagedRounds rounds in flight at once -- and the state is read from shared memory.  Layout of the staging area: st32 =
// [master | m1 | m2 | steps] (n x 4 B each), st16 = [params | ema] (n x 2 B each), n = ghi - glo.
// kStagedRounds = rounds of 4 * kTrainThreads parameters whose gradient loads are in flight together (6 cover a CTA's slice of the
// default table in one L2 round trip; the data-parallel load sums dp_world rows per group and uses 2).
template <int kStagedRounds, class LoadG, class ClearG>
__device__ __forceinline__ void grid_adam_staged(const TrainArgs& a, uint32_t base, uint32_t glo, uint32_t ghi, int tid, uint32_t s, const float* lr_tab,
                                                 const uint8_t* st32, const uint8_t* st16, LoadG load_g, ClearG clear_g) {
    const uint32_t common = a.opt_step + s, n = ghi - glo;
    for (uint32_t r0 = 4 * tid; r0 < n; r0 += 4 * kStagedRounds * kTrainThreads) {
        float4 g[kStagedRounds];
#pragma unroll
        for (int u = 0; u < kStagedRounds; ++u) {
            const uint32_t r = r0 + u * 4 * kTrainThreads;
            if (r < n) g[u] = load_g(glo + r);
        }
#pragma unroll
        for (int u = 0; u < kStagedRounds; ++u) {
            const uint32_t r = r0 + u * 4 * kTrainThreads;
            if (r >= n) continue;
            GridState4 v;
            v.w = *reinterpret_cast<const float4*>(st32 + (size_t)r * 4);
            v.m1 = *reinterpret_cast<const float4*>(st32 + ((size_t)n + r) * 4);
            v.m2 = *reinterpret_cast<const float4*>(st32 + ((size_t)2 * n + r) * 4);
            v.st = *reinterpret_cast<const uint4*>(st32 + ((size_t)3 * n + r) * 4);
            v.pr = *reinterpret_cast<const uint2*>(st16 + (size_t)r * 2);
            v.em = *reinterpret_cast<const uint2*>(st16 + ((size_t)n + r) * 2);
            grid_adam4(a, base + glo + r, glo + r, g[u], v, common, s, lr_tab, clear_g);
        }
    }
}

// Grid barrier arrival, two levels: 128 CTAs adding to ONE address serialise in the L2 atomic unit (~27 cycles per arrival:
// 1.8 us for the grid, most of the 2.3 us the barrier cost in round 1).  CTAs arrive on the counter of their group of
// kBarrierFan; the last arriver of a group (it sees the group's count complete) arrives on the top counter, which is the
// one everybody polls.  All counters only ever increase (a barrier instance adds kBarrierFan to every group counter and
// n_groups to the top one), so nothing is reset.  Ordering: acq_rel on the group counter chains the releases of the group
// into the last arriver, whose release on the top counter publishes all of them to the acquiring pollers.
#ifndef NRCB_BARRIER_FLAGS
#define NRCB_BARRIER_FLAGS 0   // 0: one arrival counter (default); 1: one flag word per CTA, no atomics -- measured 3x SLOWER
#endif                         // (barrier 2.0 -> 6.8 us: every thread of every CTA spinning on ld.acquire.gpu; profiles/r02b_experiments.md)
// `target` = the value the arrival counter has once every CTA reached this barrier instance (monotonic over the handle's
// life; the host advances it by the number of CTAs per instance).
// Flag form (A/B build `barflags`): CTA `tile` publishes flag[tile] = target with ONE release store (nothing serialises in the L2 atomic unit: 128
// same-address reds cost ~27 cycles each) and the waiters read all flags in parallel, one flag per thread -- still a single L2
// round trip, unlike the two-level tree.  Flags live behind the counter (words 32 ..), values only ever increase.
__device__ __forceinline__ void grid_arrive(const TrainArgs& a, uint32_t tile, uint32_t target) {
#if NRCB_BARRIER_TREE
    uint32_t* sub = a.barrier_counter + 32u * (1u + tile / kBarrierFan);      // 128 bytes apart: separate L2 lines
    uint32_t old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(sub) : "memory");
    if ((old + 1u) % kBarrierFan == 0u)
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.barrier_counter) : "memory");
#elif NRCB_BARRIER_FLAGS
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a.barrier_counter + 32u + tile), "r"(target) : "memory");
#else
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.barrier_counter) : "memory");
#endif
}
// Called by `stride` threads with idx = 0 .. stride-1 (a warp or the whole CTA); returns once every CTA has arrived.  The
// caller synchronises its threads afterwards (bar.sync / __syncwarp), which extends the acquire to all of them.
__device__ __forceinline__ void grid_wait(const TrainArgs& a, uint32_t target, uint32_t n_ctas, uint32_t idx, uint32_t stride) {
#if NRCB_BARRIER_FLAGS && !NRCB_BARRIER_TREE
    for (uint32_t t = idx; t < n_ctas; t += stride) {
        const uint32_t* f = a.barrier_counter + 32u + t;
        uint32_t seen, spins = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
            if (++spins > NRCB_SPIN_LIMIT) asm volatile("trap;");
        } while ((int32_t)(seen - target) < 0);
    }
#else
    if (idx == 0) {
        uint32_t seen, spins = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.barrier_counter) : "memory");
            if (++spins > NRCB_SPIN_LIMIT) asm volatile("trap;");
        } while ((int32_t)(seen - target) < 0);
    }
#endif
}

#ifndef NRCB_HASH_SPLIT_ENCODE
#define NRCB_HASH_SPLIT_ENCODE 1   // Hash training encode: 8 levels per thread of a sample, pipelined gather (0: round-1 form)
#endif
#ifndef NRCB_SCATTER_AGG
#define NRCB_SCATTER_AGG 1
#endif
#ifndef NRCB_AGG_LEVELS
#define NRCB_AGG_LEVELS 8     // coarsest levels whose scatter is pre-aggregated per tile in shared memory
#endif
#ifndef NRCB_SCATTER_V2
#define NRCB_SCATTER_V2 1     // one rolled pass over all levels between two barriers, one table per coarse level (0: a table sweep + 3 barriers per coarse level, unrolled)
#endif
#ifndef NRCB_AGG_SLOTS
#define NRCB_AGG_SLOTS 1024   // slots per coarse level (a tile sends 1024 corner updates to a level)
#endif
// ---------------------------------------------------------------- the step kernel
template <int ENC, bool MULTI>   // MULTI: a.n_steps optimizer steps in this launch (separate instantiation: the loop costs the
__global__ void __launch_bounds__(kTrainThreads, 1)   // single-step Hash kernel registers in its scatter loop)
nrc_train_step_kernel(const TrainArgs a, const GridMeta gm) {
    constexpr int E = ENC == 0 ? 80 : 64;
    constexpr int NMLP = 64 * E + 4 * 4096 + 1024;
    constexpr int CH = E / 16;   // 16-byte feature chunks per encode half
    using L = TrainSmem<E>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBar);
    uint64_t* a_ready = bars;       // workers -> issuer (count kWorkerWarps)
    uint64_t* d_ready = bars + 1;   // issuer  -> workers
    uint64_t* in_full = bars + 2;   // bulk copy of the raw queries
    uint64_t* w_full = bars + 3;    // bulk copy of the weight image
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
    float* warp_loss = reinterpret_cast<float*>(bars + 5);   // [4]
    uint64_t* st_full = bars + 7;   // bulk copy of the grid optimizer-state slice (Hash)
    uint32_t* s_lvmax = reinterpret_cast<uint32_t*>(bars + 8);   // [8] Hash scatter: per coarse level, tile maximum of |dL/dy| (float bits)

    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // provably warp-uniform
    const uint32_t tile = blockIdx.x;
    const uint32_t n_tiles = gridDim.x;              // CTAs = gradient slices of phase 3 (>= data tiles)
    const uint32_t n_data = a.n_local / kTT;         // 128-sample tiles of this launch: CTAs [0, n_data) run phase 1
    const uint32_t n_arrivals = train_barrier_arrivals(a.n_local, a.wide_grid);   // what one barrier instance adds to the counter everybody polls
    const bool has_tile = tile < n_data;
    const bool in_aligned = (reinterpret_cast<uintptr_t>(a.in) & 15) == 0 && a.perm_n == 0;   // bulk copy needs a contiguous tile
    uint32_t s = 0;                 // optimizer step within this launch
    // debug stamps: a single-step launch records its only step (slot 0 = kernel start), a multi-step launch its SECOND step
    // (slot 0 = top of that step, i.e. right after step 0's optimizer phase: the in-launch steady state)
    const uint32_t stamp_step = (MULTI && a.n_steps > 1) ? 1u : 0u;
    auto stamp = [&](int slot) { if (a.timeline && tid == 0 && s == stamp_step) a.timeline[tile * 8 + slot] = clock64(); };
    if (stamp_step == 0) stamp(0);

    if (tid == 0) {
        mbar_init(a_ready, kWorkerWarps); mbar_init(d_ready, 1); mbar_init(in_full, 1); mbar_init(w_full, 1); mbar_init(st_full, 1);
        fence_mbar_init();
        for (int i = 0; i < 8; ++i) s_lvmax[i] = 0;
    }
    if (warp == kWorkerWarps) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t sW = smem_u32(smem + L::kW), sX = smem_u32(smem + L::kX), sH = smem_u32(smem + L::kH);
    const uint32_t sDl = smem_u32(smem + L::kDelta), sG = smem_u32(smem + L::kG);
    const uint32_t n_steps = MULTI ? a.n_steps : 1;

  for (s = 0; s < n_steps; ++s) {
    if (stamp_step != 0) stamp(0);
    const uint32_t first = s * a.n_local + tile * kTT;   // first sample of this tile in this step
    // Frequency, s > 0: the CTA has only ARRIVED at the previous step's second grid barrier; the raw queries and their
    // encoding do not depend on the other CTAs, so they run while that barrier is pending, and the issuer warp completes
    // the wait before it fetches the new weight image.  (Hash: the encoding reads grid entries other CTAs updated: it waits
    // at the end of the step.)
    constexpr bool kOverlapBarrier = MULTI && ENC == 0;
    if (tid == 0 && has_tile && in_aligned) {
        mbar_arrive_expect_tx(in_full, 7680);
        bulk_g2s(smem + L::kIn, a.in + (size_t)first * 15, 7680, in_full);
    }
    if (warp == kWorkerWarps && has_tile) {
        if (s > 0) {
            if (kOverlapBarrier) {      // the whole issuer warp polls (one flag per lane and round)
                grid_wait(a, a.barrier_target + (2 * (s - 1) + 1) * n_arrivals, n_tiles, lane, 32);
                __threadfence();
                __syncwarp();
            }
            // the weight image was rewritten with generic stores by every CTA in the previous step; the grid barrier
            // ordered them, this makes them visible to the bulk copy (async proxy).
            if (lane == 0) asm volatile("fence.proxy.async;" ::: "memory");
        }
        if (lane == 0) {
            mbar_arrive_expect_tx(w_full, L::kWbytes);
            bulk_g2s(smem + L::kW, a.params_img, L::kWbytes, w_full);
        }
        __syncwarp();
    }

    if (!has_tile) {
        // a slice-only CTA (short data-parallel shard): nothing to do before the grid barrier
    } else if (warp == kWorkerWarps) {
        // ======================= MMA issuer =======================
        const bool leader = elect_one();
        uint32_t ph = 0;
        auto wait_workers = [&]() { mbar_wait_uniform(a_ready, ph); ph ^= 1; tc_fence_after(); };
        const uint32_t acc = tmem_base + kAcc;
        mbar_wait(w_full, s & 1);          // weight image landed (async proxy write -> async proxy read)
        // forward
        wait_workers();
        const bool acc16 = a.acc16 != 0;
        if (leader) { mma_fwd(acc, sX, sW, 64, E / 16, acc16); umma_commit(d_ready); }
        __syncwarp();
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            wait_workers();
            if (leader) { mma_fwd(acc, sH + l * 16384, sW + L::kWin + l * 8192, 64, 4, acc16); umma_commit(d_ready); }
            __syncwarp();
        }
        wait_workers();
        if (leader) { mma_fwd(acc, sH + 4 * 16384, sW + L::kWin + L::kWhid, 16, 4, acc16); umma_commit(d_ready); }
        __syncwarp();
        // backward: dL/dy (G) ready
        wait_workers();
        if (leader) {
            mma_dgrad(acc, sG, sW + L::kWin + L::kWhid, 16, 64, acc16);     // (W_out^T g)
            umma_commit(d_ready);
            mma_wgrad(tmem_base + wg_out_col(E), sH + 4 * 16384, sG, 16);     // dW_out^T = h5^T g
        }
        __syncwarp();
#pragma unroll
        for (int l = 3; l >= 0; --l) {   // delta_(l+2) ready in buffer (3-l)&1 ; produces delta_(l+1)
            wait_workers();
            if (leader) {
                const uint32_t dl = sDl + ((3 - l) & 1) * 16384;
                mma_dgrad(acc, dl, sW + L::kWin + l * 8192, 64, 64, acc16);
                umma_commit(d_ready);
                mma_wgrad(tmem_base + wg_hid_col(E, l), dl, sH + l * 16384, 64);  // dW_h(l+1) = delta_(l+2)^T h_(l+1)
            }
            __syncwarp();
        }
        wait_workers();                   // delta_1 ready in buffer 0
        if (leader) {
            if (ENC == 1) mma_dgrad(acc, sDl, sW, 64, 32, acc16);               // dL/d(grid features)
            mma_wgrad(tmem_base + kWgIn, sDl, sX, E);                           // dW_in = delta_1^T x
            umma_commit(d_ready);
        }
        __syncwarp();
    } else {
        // ======================= workers: two threads per sample (column halves) =======================
        const int qd = warp & 3, hf = warp >> 2;
        const uint32_t row = qd * 32 + lane;
        const uint32_t tq = tmem_base + ((qd * 32u) << 16);
        const uint32_t cb = hf * 32;          // this thread's 32 accumulator columns
        const uint32_t n_total = a.n_global * 3;
        const bool acc16 = a.acc16 != 0;     // accumulator columns hold fp16 in their low halves instead of fp32
        uint32_t ph = 0;
        auto wait_mma = [&]() { mbar_wait(d_ready, ph); ph ^= 1; tc_fence_after(); };
        auto signal = [&]() { fence_proxy_async_smem(); tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(a_ready); };

        // source record of this sample: contiguous, or gathered through the per-frame permutation (fused shuffle)
        const uint32_t src_row = a.perm_n ? permute16(a.perm_base + first + row, a.perm_seed) % a.perm_n : first + row;
        float q[15];
        if (in_aligned) {
            mbar_wait(in_full, s & 1);
            const float* s_in = reinterpret_cast<const float*>(smem + L::kIn);
#pragma unroll
            for (int i = 0; i < 15; ++i) q[i] = s_in[row * 15 + i];
        } else {
#pragma unroll
            for (int i = 0; i < 15; ++i) q[i] = __ldg(a.in + (size_t)src_row * 15 + i);
        }
        float tg[3] = {0.0f, 0.0f, 0.0f};
        if (hf == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) tg[c] = __ldg(a.targets + (size_t)src_row * 3 + c);
        }
        if constexpr (ENC == 1 && NRCB_HASH_SPLIT_ENCODE) {
            // both threads of a sample gather: 8 levels each, next level's loads in flight during the current FMA chain
            const __half2* grid = reinterpret_cast<const __half2*>(a.params + NMLP);
            uint32_t e[16];
            int ids[4];
            if (hf == 0) encode_hash_split<0>(q, grid, gm, e, ids); else encode_hash_split<1>(q, grid, gm, e, ids);
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<uint4*>(row_chunk(smem + L::kX, row, ids[c])) = make_uint4(e[4 * c], e[4 * c + 1], e[4 * c + 2], e[4 * c + 3]);
        } else {
            uint32_t e[E / 4];
            if constexpr (ENC == 0) {
                if (hf == 0) encode_frequency_half<0>(q, e); else encode_frequency_half<1>(q, e);
            } else {
                const __half2* grid = reinterpret_cast<const __half2*>(a.params + NMLP);
                if (hf == 0) encode_hash_half<0>(q, grid, gm, e); else encode_hash_half<1>(q, grid, gm, e);
            }
#pragma unroll
            for (int c = 0; c < CH; ++c)
                *reinterpret_cast<uint4*>(row_chunk(smem + L::kX, row, hf * CH + c)) = make_uint4(e[4 * c], e[4 * c + 1], e[4 * c + 2], e[4 * c + 3]);
        }
        signal();
        stamp(1);
        // forward: h_(l+1) = relu(acc) -> fp16 -> smem (the backward pass takes the ReLU mask from these fp16 activations,
        // like the reference: fully_fused_mlp.cu:150-259 reads the stored forward activations)
#pragma unroll
        for (int l = 0; l < 5; ++l) {
            wait_mma();
            uint32_t v[32];
            tmem_ld32(tq + kAcc + cb, v);
            tmem_wait_ld();
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; ++i)
                w[i] = acc16 ? relu_h2(__byte_perm(v[2 * i], v[2 * i + 1], 0x5410)) : pack_relu_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<uint4*>(row_chunk(smem + L::kH + l * 16384, row, hf * 4 + c)) =
                    make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
            signal();
        }
        // output + loss (relative_l2_luminance.h:62-86) + output-ReLU transfer (common_device.h:241-249)
        wait_mma();
        stamp(2);
        if (hf == 0) {
            float loss = 0.0f;
            uint32_t v[16];
            tmem_ld16(tq + kAcc, v);
            tmem_wait_ld();
            float y[3], gr[3];
#pragma unroll
            for (int c = 0; c < 3; ++c)
                y[c] = acc16 ? fmaxf(__half2float(__ushort_as_half((unsigned short)(v[c] & 0xFFFFu))), 0.0f)
                             : __half2float(__float2half_rn(fmaxf(__uint_as_float(v[c]), 0.0f)));
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
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
            if (lane == 0) warp_loss[qd] = loss;
        }
        signal();
        // Per-tile weight-gradient partials: M = 64 accumulators keep row o in lane (o/16)*32 + o%16.  The TMEM loads are
        // warp-wide (.sync.aligned): every lane executes them, lanes 0..15 store.  Column chunks of 16 are split between the
        // two halves.  An accumulator is dumped as soon as a commit covers its MMAs -- inside the backward chain, in the
        // shadow of the next layer's MMAs; only dW_h1 and dW_in are left for the end of the tile.
        const bool owner = lane < 16;
        const uint32_t o = qd * 16 + (lane & 15);
        float* P = a.partials + (size_t)tile * NMLP;
        auto store16 = [&](const uint32_t (&v)[16], float* dst) {   // 16 consecutive floats of row o
            if (owner) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    reinterpret_cast<float4*>(dst)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                    __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
            }
        };
        auto dump_hidden = [&](int l) {                 // dW_h(l+1), l = 0..3: this thread's 32 columns, both loads in flight
            uint32_t v0[16], v1[16];
            tmem_ld16(tq + wg_hid_col(E, l) + hf * 32, v0);
            tmem_ld16(tq + wg_hid_col(E, l) + hf * 32 + 16, v1);
            tmem_wait_ld();
            float* dst = P + 64 * E + l * 4096 + o * 64 + hf * 32;
            store16(v0, dst);
            store16(v1, dst + 16);
        };
        auto dump_out = [&]() {                         // dW_out^T [64 x 16]: element (k = o, c) -> dW_out[c][k]
            if (hf == 1) {
                uint32_t v[16];
                tmem_ld16(tq + wg_out_col(E), v);
                tmem_wait_ld();
                if (owner) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) P[64 * E + 4 * 4096 + c * 64 + o] = __uint_as_float(v[c]);
                }
            }
        };
        // backward chain: delta_l = (W^T delta_(l+1)) .* [h_l > 0], rounded to fp16
#pragma unroll
        for (int l = 4; l >= 0; --l) {
            wait_mma();
            uint8_t* dbuf = smem + L::kDelta + ((4 - l) & 1) * 16384;
            uint32_t v[32], w[16];
            tmem_ld32(tq + kAcc + cb, v);
            uint4 hh[4];                       // h_(l+1), this thread's 32 columns (its own stores of the forward pass)
#pragma unroll
            for (int c = 0; c < 4; ++c) hh[c] = *reinterpret_cast<const uint4*>(row_chunk(smem + L::kH + l * 16384, row, hf * 4 + c));
            tmem_wait_ld();
            const uint32_t* hw = reinterpret_cast<const uint32_t*>(hh);
            const __half2 zero2 = __float2half2_rn(0.0f);
#pragma unroll
            for (int i = 0; i < 16; ++i)
                w[i] = (acc16 ? __byte_perm(v[2 * i], v[2 * i + 1], 0x5410) : pack_h2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]))) &
                       __hgt2_mask(*reinterpret_cast<const __half2*>(&hw[i]), zero2);
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<uint4*>(row_chunk(dbuf, row, hf * 4 + c)) = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
            signal();
            // the commit just waited for covers the weight-gradient MMAs issued before it: dW_out (l = 3), dW_h(l+2) (l <= 2)
#if !NRCB_TIMING_SKIP_DUMP    // timing experiment only: what the in-chain dumps cost
            if (l == 3) dump_out();
            else if (l <= 2) dump_hidden(l + 1);
#endif
        }
        // all MMAs done: the remaining weight-gradient accumulators (and, Hash, dL/d(grid features)) are in TMEM
        wait_mma();
        stamp(3);
        if (warp == 0 && lane == 0) a.loss_partials[tile] = (warp_loss[0] + warp_loss[1]) + (warp_loss[2] + warp_loss[3]);
#if NRCB_SCATTER_V2
        if constexpr (ENC == 1) {
            // Grid gradient scatter (grid.h:252-255,301-319), second form: ONE pass over the 16 levels between TWO barriers.
            // Both threads of a sample walk all levels; half hf takes the four corners with dz = hf.  The kAgg coarsest levels
            // are pre-aggregated per tile (a concentrated scene sends the 1024 corner updates of a coarse level to a few dozen
            // entries, which serialises global atomics in L2): every such level owns a direct-mapped table of kSlots
            // (tag, fixed-point accumulator pair) slots in the dead activation / delta area, so no level waits for another;
            // the first form paid three CTA barriers and a table sweep per coarse level.
            constexpr int kAgg = NRCB_AGG_LEVELS;
            constexpr uint32_t kSlots = NRCB_AGG_SLOTS, kEmpty = 0xFFFFFFFFu;
            constexpr int kSlotBits = kSlots == 512 ? 9 : kSlots == 1024 ? 10 : kSlots == 2048 ? 11 : 12;
            static_assert((1u << kSlotBits) == kSlots, "NRCB_AGG_SLOTS must be 512, 1024, 2048 or 4096");
            static_assert(kAgg <= 8 && kAgg * kSlots * 12 <= 7 * 16384 && (kSlots & (kSlots - 1)) == 0, "tables live in kH .. kDelta");
            uint32_t* s_tag = reinterpret_cast<uint32_t*>(smem + L::kH);                       // [kAgg][kSlots]
            int* s_acc = reinterpret_cast<int*>(smem + L::kH + kAgg * kSlots * 4);             // [kAgg][kSlots][2] fixed point
            const uint32_t wtid = warp * 32 + lane;                                            // 0..255 over the 8 worker warps
            auto workers_sync = [] { asm volatile("bar.sync 1, 256;" ::: "memory"); };
            const bool g16 = a.grid_grad_fp16 != 0;   // reference arithmetic: half products, half2 atomics, no aggregation
            auto grad_of = [&](uint32_t w) {   // dL/d(feature) of this sample, rounded to fp16 like the reference's dL_dy
                return acc16 ? __half2float(__ushort_as_half((unsigned short)(w & 0xFFFFu))) : __half2float(__float2half_rn(__uint_as_float(w)));
            };
            if (!g16) {
                uint32_t vc[16];                       // the coarse levels' gradients: per-level maximum over the tile
                tmem_ld16(tq + kAcc, vc);
                uint4* t4 = reinterpret_cast<uint4*>(s_tag);
                for (uint32_t i = wtid; i < kAgg * kSlots / 4; i += 256) t4[i] = make_uint4(kEmpty, kEmpty, kEmpty, kEmpty);
                uint4* a4 = reinterpret_cast<uint4*>(s_acc);
                for (uint32_t i = wtid; i < kAgg * kSlots / 2; i += 256) a4[i] = make_uint4(0u, 0u, 0u, 0u);
                tmem_wait_ld();
                if (hf == 0) {
#pragma unroll
                    for (int lv = 0; lv < kAgg; ++lv) {
                        const uint32_t mx = __reduce_max_sync(0xffffffffu, max(__float_as_uint(fabsf(grad_of(vc[2 * lv]))), __float_as_uint(fabsf(grad_of(vc[2 * lv + 1])))));
                        if (lane == 0) atomicMax(&s_lvmax[lv], mx);
                    }
                }
                workers_sync();
            }
            // The level loop is NOT unrolled on purpose: a CTA walks this code once per step, so an unrolled body (16 levels x
            // 4 corners x 3 paths) is paid for in instruction fetches -- ncu attributed 69 % of the stall samples of the unrolled
            // scatter to `no instruction`.  The level's two gradient columns come straight from TMEM.
            // Coarse (shared-memory) and fine (global) levels alternate: 0, 8, 1, 9, ...
#pragma unroll 1
            for (int it = 0; it < 16; ++it) {
                const int lv = (it & 1) ? 8 + (it >> 1) : (it >> 1);
                uint32_t gw[2];
                tmem_ld2(tq + kAcc + 2 * lv, gw);
                tmem_wait_ld();
                const float g0 = grad_of(gw[0]), g1 = grad_of(gw[1]);
                if (g0 == 0.0f && g1 == 0.0f) continue;
#if NRCB_TIMING_ONLY_LEVEL >= 0  // timing experiment only: the scatter of ONE level
                if (lv != NRCB_TIMING_ONLY_LEVEL) continue;
#endif
#if NRCB_TIMING_SKIP_FINE       // timing experiments only (wrong gradients): scatter without its fine / coarse levels
                if (lv >= kAgg) continue;
#endif
#if NRCB_TIMING_SKIP_COARSE
                if (lv < kAgg) continue;
#endif
                const bool agg_level = lv < kAgg && !g16;
                const uint32_t size = gm.offset[lv + 1] - gm.offset[lv], sy = gm.stride_y[lv], sz = gm.stride_z[lv];
                const bool dense = (gm.dense_mask >> lv) & 1;
                float* gl = a.grid_grad + (size_t)gm.offset[lv] * 2;
                float p[3];
                uint32_t gp[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    p[d] = fmaf(gm.scale[lv], q[d], 0.5f);
                    const float fl = floorf(p[d]);
                    gp[d] = (uint32_t)(int)fl;
                    p[d] -= fl;
                }
                const float fz = hf ? p[2] : 1.0f - p[2];
                const uint32_t cz = gp[2] + hf;
                uint32_t idx[4];
                float wgt[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    wgt[c] = (((c & 1) ? p[0] : 1.0f - p[0]) * ((c & 2) ? p[1] : 1.0f - p[1])) * fz;
                    const uint32_t cx = gp[0] + (c & 1), cy = gp[1] + (c >> 1);
                    idx[c] = (dense ? cx + cy * sy + cz * sz : cx ^ (cy * 2654435761u) ^ (cz * 805459861u)) & (size - 1);
                }
                if (g16) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const __half2 prod = __hmul2(__float2half2_rn(wgt[c]), __floats2half2_rn(g0, g1));   // half x half, rounded to half
                        uint32_t* gh = reinterpret_cast<uint32_t*>(a.grid_grad) + gm.offset[lv] + idx[c];     // half2 accumulators
                        asm volatile("red.global.add.noftz.f16x2 [%0], %1;" ::"l"(gh), "r"(*reinterpret_cast<const uint32_t*>(&prod)) : "memory");
                    }
                    continue;
                }
                float v0[4], v1[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float wh = __half2float(__float2half_rn(wgt[c]));   // (GRAD_T)weight, grid.h:254
                    v0[c] = wh * g0; v1[c] = wh * g1;
                }
                if (agg_level) {
                    const uint32_t eb = max(s_lvmax[lv] >> 23, 24u);
                    const float S = __uint_as_float((273u - eb) << 23);     // 20 bits below the level's largest |dL/dy|, 1024 terms of headroom
                    uint32_t sl[4], old[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {   // the four claims are independent: all in flight before the first add
                        // multiplicative slot hash (top bits of idx * 2^32/phi): `idx & (kSlots - 1)` drops z on the dense
                        // level 1 (idx = x + 32 y + 1024 z), z-neighbours then fight for one slot and the loser goes to global
                        // memory, where the same entry is hit by every CTA (measured: that level alone 8.5 us instead of 4.4)
                        sl[c] = lv * kSlots + ((idx[c] * 2654435769u) >> (32 - kSlotBits));
                        old[c] = atomicCAS(&s_tag[sl[c]], kEmpty, idx[c]);
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        if (old[c] == kEmpty || old[c] == idx[c]) {
                            atomicAdd(&s_acc[2 * sl[c]], __float2int_rn(v0[c] * S));
                            atomicAdd(&s_acc[2 * sl[c] + 1], __float2int_rn(v1[c] * S));
                        } else {     // slot taken by another entry of this level: straight to global memory
                            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * idx[c]), "f"(v0[c]), "f"(v1[c]) : "memory");
                        }
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * idx[c]), "f"(v0[c]), "f"(v1[c]) : "memory");
                }
            }
            if (!g16) {   // flush: one global red per distinct entry the tile touched
                workers_sync();
                const uint4* t4 = reinterpret_cast<const uint4*>(s_tag);
                for (uint32_t i = wtid; i < kAgg * kSlots / 4; i += 256) {
                    const uint4 tg4 = t4[i];
                    if ((tg4.x & tg4.y & tg4.z & tg4.w) == kEmpty) continue;
                    const uint32_t tg_[4] = {tg4.x, tg4.y, tg4.z, tg4.w};
                    const uint32_t lv = (4 * i) / kSlots;
                    const uint32_t eb = max(s_lvmax[lv] >> 23, 24u);
                    const float Sinv = __uint_as_float((eb - 19u) << 23);
                    float* gl = a.grid_grad + (size_t)gm.offset[lv] * 2;
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (tg_[u] != kEmpty) {
                            const int2 ac = *reinterpret_cast<const int2*>(&s_acc[2 * (4 * i + u)]);
                            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * tg_[u]), "f"((float)ac.x * Sinv), "f"((float)ac.y * Sinv) : "memory");
                        }
                }
            }
        }
#else
        if constexpr (ENC == 1) {
            // Grid gradient scatter (grid.h:252-255,301-319).  Both threads of a sample walk all 16 levels; half hf takes the
            // four corners with dz = hf, so the halves finish together (with levels split 8 / 8 the half that owned the
            // aggregated coarse levels took 21 us, the other 9).
            // The kAggLevels coarsest levels are pre-aggregated per tile in a shared-memory table keyed by the entry index:
            // a concentrated scene (the Cornell box spans 0.1 of the unit cube) sends all 16384 x 8 corner updates of a
            // coarse level to a few dozen entries, which serialises global atomics in L2 (33 us measured).  After the
            // aggregation every CTA issues one global red per DISTINCT entry it touched.
            constexpr uint32_t kSlots = 2048, kEmpty = 0xFFFFFFFFu;
            constexpr int kAggLevels = NRCB_SCATTER_AGG ? NRCB_AGG_LEVELS : 0;
            uint32_t* s_tag = reinterpret_cast<uint32_t*>(smem + L::kH);            // [kSlots]   (activations are dead now)
            int* s_acc = reinterpret_cast<int*>(smem + L::kH + kSlots * 4);          // [kSlots][2] fixed point
            uint32_t* s_max = reinterpret_cast<uint32_t*>(smem + L::kH + kSlots * 12); // tile maximum of |dL/dy| (float bits)
            const uint32_t wtid = warp * 32 + lane;                                   // 0..255 over the 8 worker warps
            auto workers_sync = [] { asm volatile("bar.sync 1, 256;" ::: "memory"); };
            for (uint32_t sl = wtid; sl < kSlots; sl += 256) { s_tag[sl] = kEmpty; s_acc[2 * sl] = 0; s_acc[2 * sl + 1] = 0; }
            if (wtid == 0) *s_max = 0;
            workers_sync();
            uint32_t v[32];
            tmem_ld32(tq + kAcc, v);
            tmem_wait_ld();
            // a.grid_grad_fp16 (the "reference arithmetic" option, nrc_b200_set_accumulator): every contribution is the HALF product
            // (GRAD_T)weight * grad and is added with a half2 atomic into a half2 accumulator, exactly the reference's
            // atomicAdd(__half2) (grid.h:252-255,301-319) with its swamping of small addends; no shared-memory pre-aggregation.
            const bool g16 = a.grid_grad_fp16 != 0;
            auto level = [&](auto LV) {
                constexpr int lv = decltype(LV)::value;
                constexpr bool agg_level = lv < kAggLevels;
                const bool agg = agg_level && !g16;       // uniform over the CTA: the barriers below are executed by all or none
                const float g0 = acc16 ? __half2float(__ushort_as_half((unsigned short)(v[2 * lv] & 0xFFFFu))) : __half2float(__float2half_rn(__uint_as_float(v[2 * lv])));
                const float g1 = acc16 ? __half2float(__ushort_as_half((unsigned short)(v[2 * lv + 1] & 0xFFFFu))) : __half2float(__float2half_rn(__uint_as_float(v[2 * lv + 1])));
                const uint32_t size = gm.offset[lv + 1] - gm.offset[lv], sy = gm.stride_y[lv], sz = gm.stride_z[lv];
                const bool dense = (gm.dense_mask >> lv) & 1;
                float* gl = a.grid_grad + (size_t)gm.offset[lv] * 2;
                const bool active = g0 != 0.0f || g1 != 0.0f;
                float S = 0.0f, Sinv = 0.0f;
                if (agg_level && agg) {   // common scale of this level for the whole tile
                    const uint32_t mx = __reduce_max_sync(0xffffffffu, max(__float_as_uint(fabsf(g0)), __float_as_uint(fabsf(g1))));
                    if (lane == 0 && hf == 0) atomicMax(s_max, mx);
                    workers_sync();
                    const uint32_t eb = max(*s_max >> 23, 24u);
                    S = __uint_as_float((273u - eb) << 23);
                    Sinv = __uint_as_float((eb - 19u) << 23);
                }
                float p[3];
                uint32_t gp[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    p[d] = fmaf(gm.scale[lv], q[d], 0.5f);
                    const float fl = floorf(p[d]);
                    gp[d] = (uint32_t)(int)fl;
                    p[d] -= fl;
                }
                const float fz = hf ? p[2] : 1.0f - p[2];
                const uint32_t cz = gp[2] + hf;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float wgt = (((c & 1) ? p[0] : 1.0f - p[0]) * ((c & 2) ? p[1] : 1.0f - p[1])) * fz;
                    const uint32_t cx = gp[0] + (c & 1), cy = gp[1] + (c >> 1);
                    const uint32_t idx = (dense ? cx + cy * sy + cz * sz
                                                : cx ^ (cy * 2654435761u) ^ (cz * 805459861u)) & (size - 1);
                    const float wh = __half2float(__float2half_rn(wgt));   // (GRAD_T)weight, grid.h:254
                    const float v0 = wh * g0, v1 = wh * g1;
                    if (!active) continue;
                    if (g16) {
                        const __half2 prod = __hmul2(__float2half2_rn(wgt), __floats2half2_rn(g0, g1));   // half x half, rounded to half
                        uint32_t* gh = reinterpret_cast<uint32_t*>(a.grid_grad) + gm.offset[lv] + idx;      // half2 accumulators
                        asm volatile("red.global.add.noftz.f16x2 [%0], %1;" ::"l"(gh), "r"(*reinterpret_cast<const uint32_t*>(&prod)) : "memory");
                    } else if (!agg_level || !agg) {
                        asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * idx), "f"(v0), "f"(v1) : "memory");
                    } else {
                        // CTA-level aggregation in a shared-memory table of FIXED-POINT accumulators: 32-bit integer atomics
                        // are native in shared memory (ATOMS.ADD; float / 64-bit adds compile to CAS spin loops that collapse
                        // under same-address conflicts).  Scale: 20 bits below the largest |dL/dy| of the tile at this level,
                        // 1024 terms of headroom.
                        const uint32_t sl = idx & (kSlots - 1);
                        const uint32_t old = atomicCAS(&s_tag[sl], kEmpty, idx);
                        if (old == kEmpty || old == idx) {
                            atomicAdd(&s_acc[2 * sl], __float2int_rn(v0 * S));
                            atomicAdd(&s_acc[2 * sl + 1], __float2int_rn(v1 * S));
                        } else {
                            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * idx), "f"(v0), "f"(v1) : "memory");
                        }
                    }
                }
                if (agg_level && agg) {   // flush the table of this level: one global red per distinct entry
                    workers_sync();
                    for (uint32_t sl = wtid; sl < kSlots; sl += 256) {
                        const uint32_t tg_ = s_tag[sl];
                        if (tg_ != kEmpty) {
                            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gl + 2 * tg_), "f"((float)s_acc[2 * sl] * Sinv), "f"((float)s_acc[2 * sl + 1] * Sinv) : "memory");
                            s_tag[sl] = kEmpty; s_acc[2 * sl] = 0; s_acc[2 * sl + 1] = 0;
                        }
                    }
                    if (wtid == 0) *s_max = 0;
                    workers_sync();
                }
            };
            level(std::integral_constant<int, 0>{}); level(std::integral_constant<int, 1>{});
            level(std::integral_constant<int, 2>{}); level(std::integral_constant<int, 3>{});
            level(std::integral_constant<int, 4>{}); level(std::integral_constant<int, 5>{});
            level(std::integral_constant<int, 6>{}); level(std::integral_constant<int, 7>{});
            level(std::integral_constant<int, 8>{}); level(std::integral_constant<int, 9>{});
            level(std::integral_constant<int, 10>{}); level(std::integral_constant<int, 11>{});
            level(std::integral_constant<int, 12>{}); level(std::integral_constant<int, 13>{});
            level(std::integral_constant<int, 14>{}); level(std::integral_constant<int, 15>{});
        }
#endif
        if (a.timeline && tid == 128 && s == stamp_step) a.timeline[tile * 8 + 7] = clock64();   // half 1 done scattering (debug)
        {   // the last two accumulators: dW_in and dW_h1, all loads in flight before the wait
            constexpr int kInChunks = E / 16;                 // 5 (Frequency) / 4 (Hash)
            constexpr int kSplit = (kInChunks + 1) / 2;       // half 0: chunks [0, kSplit); half 1: the rest
            const int c0 = hf == 0 ? 0 : kSplit, nc = hf == 0 ? kSplit : kInChunks - kSplit;
            uint32_t vi[kSplit][16], h0[16], h1[16];
#pragma unroll
            for (int c = 0; c < kSplit; ++c)
                if (c < nc) tmem_ld16(tq + kWgIn + (c0 + c) * 16, vi[c]);
            tmem_ld16(tq + wg_hid_col(E, 0) + hf * 32, h0);
            tmem_ld16(tq + wg_hid_col(E, 0) + hf * 32 + 16, h1);
            tmem_wait_ld();
#pragma unroll
            for (int c = 0; c < kSplit; ++c)
                if (c < nc) store16(vi[c], P + o * E + (c0 + c) * 16);
            store16(h0, P + 64 * E + o * 64 + hf * 32);
            store16(h1, P + 64 * E + o * 64 + hf * 32 + 16);
        }
    }
    // ---- phase 2: publish this tile's partials, wait for every tile of the batch
    stamp(4);
    __threadfence();
    tc_fence_before();
    __syncthreads();
    // this CTA's slice of the grid parameters in phase 3 (multiples of 8: 16-byte bulk copies of the fp16 arrays)
    const uint32_t n_grid = a.n_params - NMLP;
    const uint32_t SG = ((n_grid + n_tiles - 1) / n_tiles + 7) & ~7u;
    const uint32_t glo = min(n_grid, tile * SG), ghi = min(n_grid, glo + SG);
    const bool staged = ENC == 1 && NRCB_ADAM_STAGED && a.apply && ghi > glo && ghi - glo <= (uint32_t)L::kStageMax;
    if constexpr (ENC == 1) {
        if (staged && tid == 32) {
            // every buffer of phase 1 is dead (the __syncthreads above): the state slice lands in shared memory while the grid
            // barrier is pending.  fence.proxy.async: generic accesses of the staging area (WAR) and the generic stores that
            // wrote this state in the previous step precede the async-proxy copies.
            const uint32_t n = ghi - glo;
            const size_t i0 = (size_t)NMLP + glo;
            asm volatile("fence.proxy.async;" ::: "memory");
            mbar_arrive_expect_tx(st_full, n * 20);
            bulk_g2s(smem + L::kStage32, a.master + i0, n * 4, st_full);
            bulk_g2s(smem + L::kStage32 + (size_t)n * 4, a.m1 + i0, n * 4, st_full);
            bulk_g2s(smem + L::kStage32 + (size_t)n * 8, a.m2 + i0, n * 4, st_full);
            bulk_g2s(smem + L::kStage32 + (size_t)n * 12, a.steps + i0, n * 4, st_full);
            bulk_g2s(smem + L::kStage16, a.params + i0, n * 2, st_full);
            bulk_g2s(smem + L::kStage16 + (size_t)n * 2, a.ema + i0, n * 2, st_full);
        }
    }
    if (tid == 0) {
        grid_arrive(a, tile, a.barrier_target + 2 * s * n_arrivals);
        for (int i = 0; i < 8; ++i) s_lvmax[i] = 0;     // the scatter of this step is over: ready for the next step's maxima
    }
    // slice `tile` of the parameters belongs to this CTA in phase 3: at most one matrix parameter per thread.  Its optimizer
    // state is loaded here, while the barrier is pending (the replicas of a data-parallel job hold identical state).
    const uint32_t S = ((NMLP + n_tiles - 1) / n_tiles + 3) & ~3u;     // parameters per CTA (multiple of 4)
    const uint32_t lo = min((uint32_t)NMLP, tile * S), hi = min((uint32_t)NMLP, lo + S);
    const bool pre_ok = a.apply && hi - lo <= (uint32_t)kTrainThreads;
    ParamState pst = {};
    if (pre_ok && (uint32_t)tid < hi - lo) pst = load_param_state(a, lo + tid);
    grid_wait(a, a.barrier_target + 2 * s * n_arrivals, n_tiles, tid, kTrainThreads);
    __syncthreads();
    tc_fence_after();
    stamp(5);

    // ---- phase 3: slice c of the gradient: fixed-order sum over tiles, then Adam + EMA (or store for all-reduce)
    {
        const uint32_t ncol4 = (hi - lo) >> 2;                             // NMLP % 4 == 0 -> whole float4 columns
        float4* scratch = reinterpret_cast<float4*>(smem + L::kScratch);   // [groups][ncol4]
        const uint32_t groups = ncol4 == 0 ? 0 : (ncol4 >= (uint32_t)kTrainThreads ? 1 : min((uint32_t)kTrainThreads / ncol4, n_data));
        for (uint32_t w = tid; w < groups * ncol4; w += kTrainThreads) {   // one pass when the batch has >= 20 tiles
            const uint32_t g = w / ncol4, c4 = w - g * ncol4;
            const float* src = a.partials + lo + 4 * c4;
            float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            // up to 24 independent L2 loads in flight: the 22 tiles a thread sums for the 16384-sample batch cost ONE L2 round
            // trip (a dependent chain of ~0.4 us hits otherwise), summed in increasing tile order
            constexpr int kInFlight = 24;
            for (uint32_t t0 = g; t0 < n_data; t0 += kInFlight * groups) {
                float4 v[kInFlight];
#pragma unroll
                for (int u = 0; u < kInFlight; ++u) {
                    const uint32_t t = t0 + u * groups;
                    v[u] = t < n_data ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)t * NMLP)) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                }
#pragma unroll
                for (int u = 0; u < kInFlight; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
            }
            scratch[w] = acc;
        }
        float* lr_tab = reinterpret_cast<float*>(smem + L::kScratch + 8192);      // behind the reduction scratch (<= 4.6 KB)
        if constexpr (ENC == 1) {
            if ((uint32_t)tid < kLrTab) {
                const uint32_t common = a.opt_step + s, cur = common > (uint32_t)tid ? common - tid : 1u;
                lr_tab[tid] = adam_bias_corrected_lr(a, cur);
            }
        }
        __syncthreads();
        const float* sc = reinterpret_cast<const float*>(scratch);
        auto slice_grad = [&](uint32_t j) {     // local gradient of parameter lo + j: groups summed in fixed order
            float gsum = 0.0f;
            for (uint32_t g = 0; g < groups; ++g) gsum += sc[(g * ncol4 + (j >> 2)) * 4 + (j & 3)];
            return gsum;
        };
        const uint32_t R = a.dp_world, me = a.dp_rank;
        if (!(a.apply && R > 1)) {
            // ---- single GPU (or gradient-only for an external all-reduce)
            for (uint32_t j = tid; j < hi - lo; j += kTrainThreads) {
                const float g = slice_grad(j);
                a.grad_f32[lo + j] = g;
                if (pre_ok) adam_ema_matrix_param(a, lo + j, g, E, s, pst);
                else if (a.apply) adam_ema_param(a, lo + j, g, E, s);
            }
            if constexpr (ENC == 1) {   // grid parameters: gradient = the fp32 scatter accumulator
                const __half* gg16 = reinterpret_cast<const __half*>(a.grid_grad);
                auto load_g = [&](uint32_t j) {
                    if (a.grid_grad_fp16) {      // half2 accumulators (reference arithmetic): four halves -> four floats, exact
                        const uint2 hv = __ldcg(reinterpret_cast<const uint2*>(gg16 + j));
                        const float2 lo2 = __half22float2(*reinterpret_cast<const __half2*>(&hv.x)), hi2 = __half22float2(*reinterpret_cast<const __half2*>(&hv.y));
                        return make_float4(lo2.x, lo2.y, hi2.x, hi2.y);
                    }
                    return __ldcg(reinterpret_cast<const float4*>(a.grid_grad + j));
                };
                auto clear_g = [&](uint32_t j) {
                    if (a.grid_grad_fp16) *reinterpret_cast<uint2*>(const_cast<__half*>(gg16) + j) = make_uint2(0u, 0u);
                    else *reinterpret_cast<float4*>(a.grid_grad + j) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                };
                if (staged) {
                    mbar_wait(st_full, s & 1);
                    grid_adam_staged<6>(a, NMLP, glo, ghi, tid, s, lr_tab, smem + L::kStage32, smem + L::kStage16, load_g, clear_g);
                } else if (a.apply) grid_adam_chunk(a, NMLP, glo, ghi, tid, E, s, lr_tab, load_g, clear_g);
                else
                    for (uint32_t j = glo + tid; j < ghi; j += kTrainThreads)
                        a.grad_f32[NMLP + j] = a.grid_grad_fp16 ? __half2float(gg16[j]) : __ldcg(a.grid_grad + j);
            }
            if (tile == 0 && warp == 0) {   // Trainer::loss (trainer.h:205-207): sum of the per-element losses
                float l = 0.0f;
                for (uint32_t t = lane; t < n_data; t += 32) l += __ldcg(a.loss_partials + t);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
                if (lane == 0) a.loss_out[s] = l;
            }
        } else {
            // ---- data parallel: the all-reduce happens here, over NVLink peer memory.
            // MLP slice + loss use a low-latency protocol: every 8-byte store carries (value, sequence tag), so the
            // receiver polls the data words themselves -- no fence, no separate flag, one NVLink one-way latency.
            // The (large) grid part is pushed as plain floats, then fence + one flag per slice.
            // seq = exchange sequence number of this step (monotonic since the buffers were connected, independent of the
            // optimizer step counter, which a weight upload resets); buffers double-buffered by its parity.
            const uint32_t seq = a.dp_seq + s;
            const uint32_t par = seq & 1;
            const uint32_t LLS = (NMLP + 1 + 31) & ~31u;                        // float2 slots of the tagged region
            auto row = [&](uint32_t dst, uint32_t src) { return a.xbuf[dst] + ((size_t)par * R + src) * a.x_stride; };
            const float tag = __uint_as_float(seq);
            // A silent peer must not take the context down: after dp_timeout_ms the waiting thread raises the host-mapped
            // error word (the next API call on the handle reports it) and gives up; the step's result is then undefined.
            auto timed_out = [&](uint32_t& spins, unsigned long long& t0) {
                if ((++spins & 1023u) != 0) return false;
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) { t0 = now; return false; }
                if (now - t0 > (unsigned long long)a.dp_timeout_ms * 1000000ull || *(volatile uint32_t*)a.error_flag != 0) {
                    *(volatile uint32_t*)a.error_flag = seq;
                    __threadfence_system();
                    return true;
                }
                return false;
            };
            float g_mine[2] = {0.0f, 0.0f};                                       // <= 2 slice elements per thread
            for (uint32_t j = tid, u = 0; j < hi - lo; j += kTrainThreads, ++u) {
                const float g = slice_grad(j);
                if (u < 2) g_mine[u] = g;
                for (uint32_t r = 0; r < R; ++r)
                    if (r != me || u >= 2)
                        asm volatile("st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(reinterpret_cast<float2*>(row(r, me)) + lo + j), "f"(g), "f"(tag) : "memory");
            }
            float l_mine = 0.0f;
            if (tile == 0 && warp == 0) {
                for (uint32_t t = lane; t < n_data; t += 32) l_mine += __ldcg(a.loss_partials + t);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) l_mine += __shfl_xor_sync(0xffffffffu, l_mine, o);
                if (lane == 0)
                    for (uint32_t r = 0; r < R; ++r)
                        if (r != me) asm volatile("st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(reinterpret_cast<float2*>(row(r, me)) + NMLP), "f"(l_mine), "f"(tag) : "memory");
            }
            if constexpr (ENC == 1) {
                for (uint32_t j = glo + 4 * tid; j < ghi; j += 4 * kTrainThreads) {   // glo, ghi multiples of 4
                    const float4 g = __ldcg(reinterpret_cast<const float4*>(a.grid_grad + j));
                    *reinterpret_cast<float4*>(a.grid_grad + j) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                    for (uint32_t r = 0; r < R; ++r) *reinterpret_cast<float4*>(row(r, me) + 2 * LLS + j) = g;
                }
                __threadfence_system();
                __syncthreads();
                if ((uint32_t)tid < R && (uint32_t)tid != me) {
                    uint32_t* f = a.xflag[tid] + ((size_t)par * R + me) * kDpFlagSlots + tile;
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(seq) : "memory");
                }
            }
            auto poll = [&](const float2* p) {      // tagged word of a peer for this step
                float2 v;
                uint32_t spins = 0;
                unsigned long long t0 = 0;
                for (;;) {
                    asm volatile("ld.relaxed.sys.global.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p) : "memory");
                    if (__float_as_uint(v.y) == seq) return v.x;
                    if (timed_out(spins, t0)) return 0.0f;
                }
            };
            for (uint32_t j = tid, u = 0; j < hi - lo; j += kTrainThreads, ++u) {
                float g = 0.0f;
                for (uint32_t r = 0; r < R; ++r)     // rank order: identical sums on every rank
                    g += (r == me && u < 2) ? g_mine[u] : poll(reinterpret_cast<const float2*>(row(me, r)) + lo + j);
                a.grad_f32[lo + j] = g;
                if (pre_ok) adam_ema_matrix_param(a, lo + j, g, E, s, pst);
                else adam_ema_param(a, lo + j, g, E, s);
            }
            if (tile == 0 && tid == 0) {
                float l = 0.0f;
                for (uint32_t r = 0; r < R; ++r) l += r == me ? l_mine : poll(reinterpret_cast<const float2*>(row(me, r)) + NMLP);
                a.loss_out[s] = l;
            }
            if constexpr (ENC == 1) {
                if ((uint32_t)tid < R && (uint32_t)tid != me) {
                    const uint32_t* w = a.xflag[me] + ((size_t)par * R + tid) * kDpFlagSlots + tile;
                    uint32_t seen, spins = 0;
                    unsigned long long t0 = 0;
                    do {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(w) : "memory");
                    } while ((int32_t)(seen - seq) < 0 && !timed_out(spins, t0));
                }
                __syncthreads();
                auto load_sum = [&](uint32_t j) {   // rank order: identical sums on every rank
                    float4 g = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                    for (uint32_t r = 0; r < R; ++r) {
                        const float4 v = __ldcv(reinterpret_cast<const float4*>(row(me, r) + 2 * LLS + j));
                        g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
                    }
                    return g;
                };
                if (staged) {
                    mbar_wait(st_full, s & 1);
                    grid_adam_staged<2>(a, NMLP, glo, ghi, tid, s, lr_tab, smem + L::kStage32, smem + L::kStage16, load_sum, [](uint32_t) {});
                } else grid_adam_chunk(a, NMLP, glo, ghi, tid, E, s, lr_tab, load_sum, [](uint32_t) {});
            }
        }
    }
    __syncthreads();
    stamp(6);
    if (s + 1 < n_steps) {
        // ---- every CTA has applied its slice: the next step may read the new weights (and reuse the partials buffer)
        __threadfence();
        fence_proxy_async_smem();
        __syncthreads();
        if (tid == 0) grid_arrive(a, tile, a.barrier_target + (2 * s + 1) * n_arrivals);
        if (!(kOverlapBarrier && has_tile)) {   // otherwise the issuer warp completes the wait after the next encode has started
            grid_wait(a, a.barrier_target + (2 * s + 1) * n_arrivals, n_tiles, tid, kTrainThreads);
            __threadfence();
        }
        __syncthreads();
    }
  }
    tc_fence_before();
    __syncthreads();
    if (warp == kWorkerWarps) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- stand-alone optimizer (data-parallel path)
// After the all-reduce of grad_f32: the same per-parameter update on every rank; clears the grid accumulator.
__global__ void __launch_bounds__(256)
nrc_adam_ema_kernel(const TrainArgs a, uint32_t E) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_params) return;
    adam_ema_param(a, i, a.grad_f32[i], E);
    if (i >= a.n_mlp) a.grid_grad[i - a.n_mlp] = 0.0f;
}

__global__ void __launch_bounds__(256)
nrc_make_image_kernel(const __half* __restrict__ params, __half* __restrict__ img, uint32_t E, uint32_t n_mlp) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_mlp) img[img_index(i, E)] = params[i];
}

// ---------------------------------------------------------------- launchers
cudaError_t configure_train_kernels() {   // per device (the attribute is not per process): called when a handle is created
    cudaError_t e = cudaFuncSetAttribute(nrc_train_step_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrainSmem<80>::kTotal);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(nrc_train_step_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrainSmem<80>::kTotal);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(nrc_train_step_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrainSmem<64>::kTotal);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(nrc_train_step_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrainSmem<64>::kTotal);
    return e;
}
template <int ENC>
static cudaError_t launch_train_t(const TrainArgs& a, const GridMeta& gm, cudaStream_t stream) {
    using L = TrainSmem<ENC == 0 ? 80 : 64>;
    // Cooperative launch: the grid barrier needs every CTA resident (<= 128 CTAs, one per SM).
    void* args[2] = {const_cast<TrainArgs*>(&a), const_cast<GridMeta*>(&gm)};
    const bool multi = a.apply && a.n_steps > 1;
    const void* fn = multi ? reinterpret_cast<const void*>(&nrc_train_step_kernel<ENC, true>)
                           : reinterpret_cast<const void*>(&nrc_train_step_kernel<ENC, false>);
    return cudaLaunchCooperativeKernel(fn, dim3(train_grid(a.n_local, a.wide_grid)), dim3(kTrainThreads), args, L::kTotal, stream);
}
cudaError_t launch_train_step(int encoding, const TrainArgs& a, const GridMeta& gm, cudaStream_t stream) {
    return encoding == 0 ? launch_train_t<0>(a, gm, stream) : launch_train_t<1>(a, gm, stream);
}
cudaError_t launch_adam_ema(const TrainArgs& a, cudaStream_t stream) {
    const uint32_t E = a.n_mlp == 64 * 80 + 4 * 4096 + 1024 ? 80 : 64;
    nrc_adam_ema_kernel<<<(a.n_params + 255) / 256, 256, 0, stream>>>(a, E);
    return cudaGetLastError();
}
cudaError_t launch_make_image(const __half* params, __half* params_img, uint32_t enc_width, uint32_t n_mlp,
                              cudaStream_t stream) {
    nrc_make_image_kernel<<<(n_mlp + 255) / 256, 256, 0, stream>>>(params, params_img, enc_width, n_mlp);
    return cudaGetLastError();
}

}  // namespace nrcb
