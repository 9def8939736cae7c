// Fused NRC cache query for sm_100a: raw 60-byte RadianceQuery records in, float3 radiance out.
// Replaces the reference chain  CompositeEncoding::forward_impl -> kernel_mlp_fused<64,...,INFERENCE> ->
// trim_and_cast  (tcnn/src/fully_fused_mlp.cu:499-557, nrc/src/NRCNetwork.cu:64-83)
// with ONE persistent kernel: encoded features and activations never touch HBM.
//
// CTA = 5 groups x 4 warps (640 threads). Each group owns one 128-query tile at a time; five tiles are in flight per SM
// (TMEM: 64 fp32 accumulator + 32 operand columns per tile) so that the tensor pipe, the TMEM ports and the FP32 pipes
// overlap across groups:
//   all warps : bulk-copy the tile's 7680 input bytes to smem, encode in registers, write the fp16 features as the
//               K-major A image of the input layer to shared memory (80 features would need 40 TMEM columns and cap
//               the SM at four tiles); then per layer read the fp32 accumulator from TMEM, ReLU, round to fp16 and
//               write it back to TMEM as the next layer's A operand (the activations never leave the SM).
//   warp 0 of the group additionally issues the layer: after a named barrier tells it that the A operand is complete,
//               one elected lane issues tcgen05.mma.kind::f16 (M=128 queries, N=64|16, K=16; A from shared memory for
//               the input layer, from TMEM for the others; B = the layer's weight matrix, resident in smem for the
//               whole kernel), commits to an mbarrier, waits for it and releases the other three warps through a
//               second named barrier (only one warp per group polls; the others block in bar.sync, which costs no
//               issue slots -- with dedicated issuer warps and every worker polling, 28 % of all issued instructions
//               were polls; profiles/r01_infer_experiments.md).
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <type_traits>
#include "nrc_kernels.h"
#include "nrc_encode.cuh"

#ifndef NRCB_INFER_DIRECT_WAIT
#define NRCB_INFER_DIRECT_WAIT 0      // 1: every warp of a group waits on the commit mbarrier itself -- measured slower (164.3 vs 162.0 us, A/B build directwait)
#endif
#ifndef NRCB_TIMING_SKIP_ENCODE
#define NRCB_TIMING_SKIP_ENCODE 0     // timing experiment only (wrong radiance): Frequency without the encode slices of the next tile, Hash without the gathers
#endif
namespace nrcb {

#ifndef NRC_HASH_GATHER
#define NRC_HASH_GATHER 1         // 0: level after level with 8 loads each (round-1 form, A/B builds); 1: pipelined gather (nrc_encode.cuh)
#endif
#ifndef NRC_HASH_DEPTH
#define NRC_HASH_DEPTH 2          // levels of the hash gather whose loads are in flight per thread.  Measured (A/B builds hashd3 / hashd4,
                                  // which read the twelve non-position inputs only after the gather to pay for the registers): 529 / 540 /
                                  // 531 us per 2K frame for 2 / 3 / 4 -- the gather is bound by the L1TEX tag rate (8 cycles per divergent
                                  // LDG), not by latency, once two levels overlap
#endif
#ifndef NRC_HASH_SMEM0
#define NRC_HASH_SMEM0 0          // 1: grid level 0 staged in shared memory (A/B builds: slower, it takes L1 capacity from the fine levels)
#endif
#ifndef NRC_INFER_PIPELINE
#define NRC_INFER_PIPELINE 1      // 0: encode, then the six layers, tile after tile (A/B builds; the Hash kernel always runs that loop)
#endif
constexpr int kGroups = 5;
constexpr int kTile = 128;
constexpr int kThreads = kGroups * 128;
constexpr uint32_t kTmemGroupCols = 96;   // per group: D at +0 (64 columns), A at +64 (32 columns)
constexpr uint32_t kTmemA = 64;

template <int E>
struct InferSmem {
    static constexpr int kWin = 64 * E * 2;           // W_in image
    static constexpr int kWhid = 4 * 64 * 64 * 2;     // 4 hidden images
    static constexpr int kWout = 16 * 64 * 2;
    static constexpr int kWeights = kWin + kWhid + kWout;
    static constexpr int kStage = kTile * 15 * 4;     // 7680 B of raw queries per group
    static constexpr int kOut = kTile * 3 * 4;        // 1536 B of results per group
    // Frequency only (E = 80): encoded tile = K-major A image of the input layer, per group.  Hash features (E = 64) fit
    // the 32 operand columns in TMEM, and its gathers want the shared-memory carve-out small (L1 capacity).
    static constexpr int kA0 = E > 64 ? kTile * E * 2 : 0;
    // Hash only: level 0 of the grid (16^3 = 4096 entries x 2 fp16 = 16 KB, grid.h:692-723) staged in shared memory for the whole
    // kernel: its lookups are the least coherent per byte of table (every query of an incoherent batch hits it) and an LDS costs
    // ~1/8 of the L1TEX wavefronts of a divergent LDG.  The finer levels (128 KB each and up) stay in L1 / L2.
    static constexpr int kL0 = (E > 64 || !NRC_HASH_SMEM0 || !NRC_HASH_GATHER) ? 0 : 4096 * 4;
    static constexpr int kL0Off = kWeights;
    static constexpr int kA0Off = kL0Off + kL0;
    static constexpr int kInOff = kA0Off + kGroups * kA0;
    static constexpr int kOutOff = kInOff + kGroups * kStage;
    static constexpr int kAcc = kTile * 12 + kTile * 16;   // fused accumulation: throughput (1536 B) + output (2048 B) of a tile
    static constexpr int kBarOff = kOutOff + kGroups * kOut;
    static constexpr int kAccOff = kBarOff + 256;                      // [kGroups][2 parities][kAcc], only with accumulation
    static constexpr int kTotalNoAcc = kAccOff;                        // launches without accumulation leave more L1
    static constexpr int kTotal = kAccOff + kGroups * 2 * kAcc;
};

// Named barriers of group g (128 threads each): 1 + g "A operand ready", 1 + kGroups + g "accumulator ready" / group sync.
__device__ __forceinline__ void bar_sync_id(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void bar_arrive_id(int id) { asm volatile("bar.arrive %0, 128;" ::"r"(id) : "memory"); }

// One layer of the group's tile, issued by a single elected lane.  LY is a compile-time constant so every descriptor is
// a uniform-register immediate offset and ptxas emits back-to-back UTCHMMA.
template <int ENC, bool ACC16, int LY>
__device__ __forceinline__ void issue_layer(uint32_t sbase, uint32_t a0, uint32_t dD, uint64_t* d_ready) {
    constexpr int E = ENC == 0 ? 80 : 64;
    using L = InferSmem<E>;
    constexpr uint32_t idesc64 = ACC16 ? umma_idesc_f16_acc16(128, 64, 0, 0) : umma_idesc_f16(128, 64, 0, 0);
    constexpr uint32_t idesc16 = ACC16 ? umma_idesc_f16_acc16(128, 16, 0, 0) : umma_idesc_f16(128, 16, 0, 0);
    const uint32_t dA = dD + kTmemA;
    if constexpr (LY == 0) {
        const uint64_t d_a0 = umma_desc(a0, 2048, 128);   // K-major A [128 x E]: +4096 B per k-step
        const uint64_t d_in = umma_desc(sbase, 64 * 16, 128);
#pragma unroll
        for (int k = 0; k < E / 16; ++k) {
            if constexpr (L::kA0 != 0) umma_ss(dD, d_a0 + (uint64_t)(k * 256), d_in + (uint64_t)(k * 128), idesc64, k > 0);
            else umma_ts(dD, dA + 8 * k, d_in + (uint64_t)(k * 128), idesc64, k > 0);
        }
    } else if constexpr (LY < 5) {
        const uint64_t d_hid = umma_desc(sbase + L::kWin + (LY - 1) * 8192, 64 * 16, 128);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ts(dD, dA + 8 * k, d_hid + (uint64_t)(k * 128), idesc64, k > 0);
    } else {
        const uint64_t d_out = umma_desc(sbase + L::kWin + L::kWhid, 16 * 16, 128);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ts(dD, dA + 8 * k, d_out + (uint64_t)(k * 32), idesc16, k > 0);
    }
    umma_commit(d_ready);
}

// Hash encode of one query straight into the TMEM A operand (32 words = 64 fp16 features at columns tA .. tA+31): the 16 grid
// levels are software-pipelined (level l+1's loads in flight during level l's FMA chain, nrc_encode.cuh) and leave for TMEM in
// chunks of 8 words, so at most 8 result words + two levels of pending loads are live.
template <int L, int D>
__device__ __forceinline__ void hash_levels_pipelined(GridFetch (&F)[D], uint32_t (&w)[8], const float (&q)[15], const __half2* grid,
                                                      uint32_t smem_l0, const GridMeta& gm, uint32_t tA) {
    if constexpr (L < 16) {
        if constexpr (L + D - 1 < 16) grid_issue_level<L + D - 1, NRC_HASH_SMEM0 != 0>(F[(L + D - 1) % D], q, grid, smem_l0, gm);
        w[L & 7] = grid_consume(F[L % D]);
        if constexpr ((L & 7) == 7) tmem_st8(tA + (L >> 3) * 8, w);
        hash_levels_pipelined<L + 1, D>(F, w, q, grid, smem_l0, gm, tA);
    }
}
template <int L, int D>
__device__ __forceinline__ void hash_prologue(GridFetch (&F)[D], const float (&q)[15], const __half2* grid, uint32_t smem_l0, const GridMeta& gm) {
    if constexpr (L < D - 1) {
        grid_issue_level<L, NRC_HASH_SMEM0 != 0>(F[L], q, grid, smem_l0, gm);
        hash_prologue<L + 1, D>(F, q, grid, smem_l0, gm);
    }
}
// Grid part (words 0..15): needs q[0..2] only.  NRC_HASH_DEPTH levels of loads are in flight per thread.
__device__ __forceinline__ void encode_hash_grid_to_tmem(const float (&q)[15], const __half2* grid, uint32_t smem_l0, const GridMeta& gm, uint32_t tA) {
    GridFetch F[NRC_HASH_DEPTH];
    uint32_t w[8];
    hash_prologue<0, NRC_HASH_DEPTH>(F, q, grid, smem_l0, gm);
    hash_levels_pipelined<0, NRC_HASH_DEPTH>(F, w, q, grid, smem_l0, gm, tA);
}
// One-blob, albedos, pad (words 16..31): needs q[3..14].
__device__ __forceinline__ void encode_hash_rest_to_tmem(const float (&q)[15], uint32_t tA) {
    uint32_t ob[12];
#pragma unroll
    for (int d = 0; d < 6; d += 2) oneblob_pair<false>(q[3 + d], q[4 + d], ob + 2 * d);
    {
        const uint32_t v[8] = {ob[0], ob[1], ob[2], ob[3], ob[4], ob[5], ob[6], ob[7]};
        tmem_st8(tA + 16, v);
    }
    const uint32_t v[8] = {ob[8], ob[9], ob[10], ob[11], pack_h2(q[9], q[10]), pack_h2(q[11], q[12]), pack_h2(q[13], q[14]), 0x3C003C00u};
    tmem_st8(tA + 24, v);
}

// ENC: 0 = Frequency (E = 80), 1 = Hash (E = 64).  ACC16: fp16 accumulators (one fp16 per 32-bit TMEM column, rounded after
// every K = 16 instruction: the arithmetic of the reference's wmma::accumulator<__half>, fully_fused_mlp.cu:68) instead of fp32.
template <int ENC, bool ACC16>
__global__ void __launch_bounds__(kThreads, 1)
nrc_infer_kernel(const float* __restrict__ in, float* __restrict__ out, uint32_t n, const __half* __restrict__ params,
                 const __half* __restrict__ w_img, GridMeta gm, AccumArgs acc) {
    constexpr int E = ENC == 0 ? 80 : 64;
    using L = InferSmem<E>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    uint64_t* in_full = bars;              // [kGroups] raw tile landed in smem
    uint64_t* d_ready = bars + kGroups;    // [kGroups] accumulator complete
    uint64_t* acc_full = bars + 2 * kGroups;   // [kGroups] throughput + output rows of the tile landed (pipelined loop)
    uint64_t* w_full = bars + 3 * kGroups;     // the weight image landed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * kGroups + 1);

    // warp index through a shuffle: the compiler then knows it is warp-uniform and keeps everything derived from it (TMEM
    // addresses, barrier ids, shared-memory bases of the group) in uniform registers instead of R2UR moves per use
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const uint32_t n_tiles = (n + kTile - 1) / kTile;
    const uint32_t tile_stride = gridDim.x * kGroups;

    // ---- one-time setup: barriers, TMEM, and ONE bulk copy (UBLKCP) of the weights: w_img is the UMMA K-major image of the
    //      six matrices (off(o,k) = (k/8)*(n_out*16) + (o/8)*128 + (o%8)*16 + (k%8)*2 per matrix), kept current in global
    //      memory by the optimizer phase of the training kernel, so no CTA re-formats the 45 KB with generic loads.
    if (tid == 0) {
        for (int g = 0; g < kGroups; ++g) { mbar_init(&in_full[g], 1); mbar_init(&d_ready[g], 1); mbar_init(&acc_full[g], 1); }
        mbar_init(w_full, 1);
        fence_mbar_init();
        mbar_arrive_expect_tx(w_full, L::kWeights + L::kL0);
        bulk_g2s(smem, w_img, L::kWeights, w_full);
        if constexpr (L::kL0 != 0) bulk_g2s(smem + L::kL0Off, params + (64 * E + 4 * 4096 + 16 * 64), L::kL0, w_full);   // grid level 0
    }
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    {
        const int g = warp >> 2;
        const int gtid = tid & 127;                     // 0..127 within the group
        const uint32_t row = (warp & 3) * 32 + lane;    // TMEM lane == query row inside the tile
        const uint32_t dD = tmem_base + g * kTmemGroupCols;             // accumulator, lane 0 (MMA addresses)
        const uint32_t tD = dD + (((warp & 3) * 32u) << 16);            // this warp's lane quadrant
        const uint32_t tA = tD + kTmemA;
        const uint32_t a0 = smem_u32(smem + L::kA0Off + g * L::kA0);
        uint8_t* a_row = smem + L::kA0Off + g * L::kA0 + (row >> 3) * 128 + (row & 7) * 16;   // chunk c of this row at + c*2048
        float* s_in = reinterpret_cast<float*>(smem + L::kInOff + g * L::kStage);
        float* s_out = reinterpret_cast<float*>(smem + L::kOutOff + g * L::kOut);
        uint8_t* s_acc = smem + L::kAccOff + g * 2 * L::kAcc;
        const uint32_t sbase = smem_u32(smem);
        const __half2* grid = reinterpret_cast<const __half2*>(params + (64 * E + 4 * 4096 + 16 * 64));
        const bool in_bulk_ok = (reinterpret_cast<uintptr_t>(in) & 15) == 0;
        const bool out_vec_ok = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
        const bool acc_bulk_ok = acc.out != nullptr && (reinterpret_cast<uintptr_t>(acc.throughput) & 15) == 0 &&
                                 (reinterpret_cast<uintptr_t>(acc.out) & 15) == 0;
        // tiles whose 128 queries are all pixels prefetch throughput + output rows with the same mbarrier as the input
        auto acc_staged = [&](uint32_t t) { return acc_bulk_ok && in_bulk_ok && (t + 1) * kTile <= acc.n && (t + 1) * kTile <= n; };
        const int bar_a = 1 + g, bar_d = 1 + kGroups + g;

        auto load_tile = [&](uint32_t t, uint32_t par) {  // called by all 128 threads of the group, after a group sync
            const uint32_t rows = min((uint32_t)kTile, n - t * kTile);
            const float* src = in + (size_t)t * kTile * 15;
            if (in_bulk_ok && rows == kTile) {
                if (gtid == 0) {
                    const bool st = acc_staged(t);
                    mbar_arrive_expect_tx(&in_full[g], L::kStage + (st ? L::kAcc : 0));
                    bulk_g2s_stream(s_in, src, L::kStage, &in_full[g]);
                    if (st) {
                        bulk_g2s_stream(s_acc + par * L::kAcc, acc.throughput + (size_t)t * kTile * 3, kTile * 12, &in_full[g]);
                        bulk_g2s(s_acc + par * L::kAcc + kTile * 12, acc.out + (size_t)t * kTile, kTile * 16, &in_full[g]);
                    }
                }
            } else {  // ragged tail / unaligned caller buffer: plain loads, zero-fill the missing rows
                for (uint32_t i = gtid; i < kTile * 15; i += 128) s_in[i] = i < rows * 15 ? __ldg(src + i) : 0.0f;
                bar_sync_id(bar_d);
                if (gtid == 0) mbar_arrive(&in_full[g]);
            }
        };

        uint32_t ph_d = 0;
        // The A operand of layer LY is complete (shared memory for the input layer, TMEM otherwise) and the accumulator is
        // drained: run the layer, return when its accumulator is complete.
        // `overlap` is work of this thread that does not touch the tile in flight (a slice of the NEXT tile's encoding): it
        // runs between the hand-off to the tensor pipe and the wait for the accumulator.
        auto run_layer = [&](auto LY, auto&& overlap) {
            if (decltype(LY)::value == 0 && L::kA0 != 0) fence_proxy_async_smem();   // feature image: generic writes -> tensor pipe
            else tmem_wait_st();
            tc_fence_before();
            if ((warp & 3) == 0) {
                bar_sync_id(bar_a);
                tc_fence_after();
                if (elect_one()) issue_layer<ENC, ACC16, decltype(LY)::value>(sbase, a0, dD, &d_ready[g]);
                __syncwarp();
                overlap();
                mbar_wait(&d_ready[g], ph_d);
#if !NRCB_INFER_DIRECT_WAIT
                bar_sync_id(bar_d);
#endif
            } else {
                bar_arrive_id(bar_a);
                overlap();
#if NRCB_INFER_DIRECT_WAIT
                // every warp of the group waits for the commit itself (the next commit cannot happen before all four warps have
                // passed this wait and arrived at bar_a again, so no warp can miss a phase): one hop less than being released by
                // the issuing warp through a second named barrier
                mbar_wait(&d_ready[g], ph_d);
#else
                bar_sync_id(bar_d);
#endif
            }
            ph_d ^= 1;
            tc_fence_after();
        };
        auto nothing = [] {};
        auto hidden_epilogue = [&]() {   // D (fp32) -> relu -> fp16 -> A
            // 16-column pieces, the next piece's TMEM load in flight while the current one is converted (tcgen05.wait::ld waits
            // for every outstanding load, so a load is issued right AFTER the wait that completes its predecessor)
            uint32_t v[2][16], w[8];
            tmem_ld16(tD, v[0]);
#pragma unroll
            for (int pc = 0; pc < 4; ++pc) {
                tmem_wait_ld();
                if (pc < 3) tmem_ld16(tD + (pc + 1) * 16, v[(pc + 1) & 1]);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if constexpr (ACC16) w[i] = relu_h2(__byte_perm(v[pc & 1][2 * i], v[pc & 1][2 * i + 1], 0x5410));   // fp16 in the low halves
                    else w[i] = pack_relu_h2(__uint_as_float(v[pc & 1][2 * i]), __uint_as_float(v[pc & 1][2 * i + 1]));
                }
                tmem_st8(tA + pc * 8, w);
            }
        };

        // ---- output layer done: relu, fp16 rounding (the reference's output is half), first 3 of 16 lanes; optional fused
        //      accumulate_render_radiance (nrc_helpers.cu:77-129) for the pixel queries -- the radiance is still in registers,
        //      so the separate pass (re-read 12 B radiance, 12 B throughput, 16 B output per pixel) disappears.
        //      `staged`: this tile's throughput / output rows in shared memory, or nullptr (read them from global memory).
        auto drain_output = [&](float (&y)[3]) {
            uint32_t v[4];
            tmem_ld4(tD, v);
            tmem_wait_ld();
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                if constexpr (ACC16) y[c] = fmaxf(__half2float(__ushort_as_half((unsigned short)(v[c] & 0xFFFFu))), 0.0f);
                else y[c] = __half2float(__float2half_rn(fmaxf(__uint_as_float(v[c]), 0.0f)));
            }
            // D is drained: the next tile's first run_layer orders this read before the MMA that overwrites it.
        };
        auto emit_tile = [&](uint32_t t, const uint8_t* staged, const float (&y)[3]) {
            const uint32_t rows = min((uint32_t)kTile, n - t * kTile);
            if (acc.out != nullptr) {
                const uint32_t px = t * kTile + row;
                if (px < acc.n && row < rows) {
                    float t0, t1, t2;
                    float4 d;
                    if (staged) {
                        const float* st = reinterpret_cast<const float*>(staged);
                        t0 = st[3 * row]; t1 = st[3 * row + 1]; t2 = st[3 * row + 2];
                        d = reinterpret_cast<const float4*>(staged + kTile * 12)[row];
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
                if (staged) fence_proxy_async_smem();            // order these generic reads before the next bulk write
            }
            float* dst = out + (size_t)t * kTile * 3;
            if (out_vec_ok && rows == kTile) {
                s_out[row * 3 + 0] = y[0]; s_out[row * 3 + 1] = y[1]; s_out[row * 3 + 2] = y[2];
                bar_sync_id(bar_d);
                if (gtid < 96) st_stream_f4(reinterpret_cast<float4*>(dst) + gtid, reinterpret_cast<const float4*>(s_out)[gtid]);
            } else if (row < rows) {
                dst[row * 3 + 0] = y[0]; dst[row * 3 + 1] = y[1]; dst[row * 3 + 2] = y[2];
            }
        };
        auto finish_tile = [&](uint32_t t, const uint8_t* staged) {
            float y[3];
            drain_output(y);
            emit_tile(t, staged, y);
        };

        // tile t of round r belongs to group g of CTA b with t = r*stride + g*gridDim + b: the tiles of the last (partial) round
        // land on DIFFERENT SMs (one group each, which then has the SM to itself) instead of filling a few SMs five deep
        uint32_t t = g * gridDim.x + blockIdx.x;
        uint32_t ph_in = 0;
        if constexpr (NRC_INFER_PIPELINE && L::kA0 != 0) {
            // ---- Frequency: the encoding of tile t + stride is cut into five slices that run while the tensor pipe works on
            //      layers 1..5 of tile t (the features go to the shared-memory A image, which the input layer of tile t has
            //      finished reading by then).  The raw queries of the next tile wait in registers; the stage is refilled two
            //      tiles ahead; throughput / output rows of a tile are prefetched at its start on their own mbarrier.
            auto load_in = [&](uint32_t tl) {      // all 128 threads of the group, after a group sync that freed the stage
                const uint32_t rows = min((uint32_t)kTile, n - tl * kTile);
                const float* src = in + (size_t)tl * kTile * 15;
                if (in_bulk_ok && rows == kTile) {
                    if (gtid == 0) {
                        mbar_arrive_expect_tx(&in_full[g], L::kStage);
                        bulk_g2s_stream(s_in, src, L::kStage, &in_full[g]);
                    }
                } else {  // ragged tail / unaligned caller buffer: plain loads, zero-fill the missing rows
                    for (uint32_t i = gtid; i < kTile * 15; i += 128) s_in[i] = i < rows * 15 ? __ldg(src + i) : 0.0f;
                    bar_sync_id(bar_d);
                    if (gtid == 0) mbar_arrive(&in_full[g]);
                }
            };
            float q[15];
            auto read_q = [&]() {
                mbar_wait(&in_full[g], ph_in); ph_in ^= 1;
#pragma unroll
                for (int i = 0; i < 15; ++i) q[i] = s_in[row * 15 + i];
                fence_proxy_async_smem();                    // order these generic reads before the next bulk write
            };
            uint32_t ph_acc = 0;
            if (t < n_tiles) {
                load_in(t);
                read_q();
                bar_sync_id(bar_d);                          // everyone holds its query: the stage is free
                if (t + tile_stride < n_tiles) load_in(t + tile_stride);
                encode_frequency_slice<0>(q, a_row); encode_frequency_slice<1>(q, a_row); encode_frequency_slice<2>(q, a_row);
                encode_frequency_slice<3>(q, a_row); encode_frequency_slice<4>(q, a_row);
            }
            if ((warp & 3) == 0) mbar_wait(w_full, 0);   // issuing warp: the weight image has landed (async-proxy write -> async-proxy read)
            for (; t < n_tiles; t += tile_stride) {
                const bool has_next = t + tile_stride < n_tiles;
                run_layer(std::integral_constant<int, 0>{}, [&] { if (has_next) read_q(); });   // raw queries of the next tile -> registers
                if (t + 2 * tile_stride < n_tiles) load_in(t + 2 * tile_stride);
                const bool st = acc_staged(t);
                if (st && gtid == 0) {
                    mbar_arrive_expect_tx(&acc_full[g], L::kAcc);
                    bulk_g2s_stream(s_acc, acc.throughput + (size_t)t * kTile * 3, kTile * 12, &acc_full[g]);
                    bulk_g2s(s_acc + kTile * 12, acc.out + (size_t)t * kTile, kTile * 16, &acc_full[g]);
                }
                hidden_epilogue();
                run_layer(std::integral_constant<int, 1>{}, [&] { if (has_next && !NRCB_TIMING_SKIP_ENCODE) encode_frequency_slice<0>(q, a_row); }); hidden_epilogue();
                run_layer(std::integral_constant<int, 2>{}, [&] { if (has_next && !NRCB_TIMING_SKIP_ENCODE) encode_frequency_slice<1>(q, a_row); }); hidden_epilogue();
                run_layer(std::integral_constant<int, 3>{}, [&] { if (has_next && !NRCB_TIMING_SKIP_ENCODE) encode_frequency_slice<2>(q, a_row); }); hidden_epilogue();
                run_layer(std::integral_constant<int, 4>{}, [&] { if (has_next && !NRCB_TIMING_SKIP_ENCODE) encode_frequency_slice<3>(q, a_row); }); hidden_epilogue();
                run_layer(std::integral_constant<int, 5>{}, [&] { if (has_next && !NRCB_TIMING_SKIP_ENCODE) encode_frequency_slice<4>(q, a_row); });
                if (st) { mbar_wait(&acc_full[g], ph_acc); ph_acc ^= 1; }
                finish_tile(t, st ? s_acc : nullptr);
            }
        } else {
        uint32_t par = 0;
        if (t < n_tiles) load_tile(t, 0);
        if (ENC == 1 || (warp & 3) == 0) mbar_wait(w_full, 0);   // weight image (issuing warp) and staged grid level 0 (every gathering thread) have landed
        for (; t < n_tiles; t += tile_stride) {
            mbar_wait(&in_full[g], ph_in); ph_in ^= 1;
            float q[15];
            if constexpr (ENC == 1 && NRC_HASH_GATHER != 0 && NRC_HASH_DEPTH > 2) {
                // deeper pipelines: only the position is needed for the gather; the other twelve inputs are read from the stage
                // AFTER it, so their registers carry more pending loads instead (the stage is released, and the next tile's bulk
                // copy issued, only then: it still has the six layers of this tile to land)
#pragma unroll
                for (int i = 0; i < 3; ++i) q[i] = s_in[row * 15 + i];
                encode_hash_grid_to_tmem(q, grid, smem_u32(smem + L::kL0Off), gm, tA);
#pragma unroll
                for (int i = 3; i < 15; ++i) q[i] = s_in[row * 15 + i];
                fence_proxy_async_smem();                    // order these generic reads before the next bulk write
                bar_sync_id(bar_d);                          // everyone holds its query: the stage is free
                if (t + tile_stride < n_tiles) load_tile(t + tile_stride, par ^ 1);
                encode_hash_rest_to_tmem(q, tA);
            } else {
#pragma unroll
            for (int i = 0; i < 15; ++i) q[i] = s_in[row * 15 + i];
            fence_proxy_async_smem();                        // order these generic reads before the next bulk write
            bar_sync_id(bar_d);                              // everyone holds its query: the stage is free
            if (t + tile_stride < n_tiles) load_tile(t + tile_stride, par ^ 1);
            }

            // ---- encode -> A operand of the input layer (its previous contents were consumed long ago)
            if constexpr (ENC == 1 && NRC_HASH_GATHER != 0) {
                if constexpr (NRC_HASH_DEPTH <= 2) {
                    if (!NRCB_TIMING_SKIP_ENCODE)      // timing experiment only: the query kernel without its hash-grid gathers
                    encode_hash_grid_to_tmem(q, grid, smem_u32(smem + L::kL0Off), gm, tA);
                    encode_hash_rest_to_tmem(q, tA);
                }
            } else {
                uint32_t e[E / 2];
                if constexpr (ENC == 0) encode_frequency(q, e);
                else encode_hash(q, grid, gm, e);
                if constexpr (L::kA0 != 0) {
#pragma unroll
                    for (int c = 0; c < E / 8; ++c)
                        *reinterpret_cast<uint4*>(a_row + c * 2048) = make_uint4(e[4 * c], e[4 * c + 1], e[4 * c + 2], e[4 * c + 3]);
                } else {
#pragma unroll
                    for (int c = 0; c < E / 16; ++c) {
                        uint32_t v[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) v[i] = e[c * 8 + i];
                        tmem_st8(tA + c * 8, v);
                    }
                }
            }
            run_layer(std::integral_constant<int, 0>{}, nothing); hidden_epilogue();
            run_layer(std::integral_constant<int, 1>{}, nothing); hidden_epilogue();
            run_layer(std::integral_constant<int, 2>{}, nothing); hidden_epilogue();
            run_layer(std::integral_constant<int, 3>{}, nothing); hidden_epilogue();
            run_layer(std::integral_constant<int, 4>{}, nothing); hidden_epilogue();
            run_layer(std::integral_constant<int, 5>{}, nothing);
            finish_tile(t, acc_staged(t) ? s_acc + par * L::kAcc : nullptr);
            par ^= 1;
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 512);
}

template <int ENC, bool ACC16>
static cudaError_t launch_infer_t(const float* in, float* out, uint32_t n, const __half* params, const __half* w_img, const GridMeta& gm,
                                  int num_sms, cudaStream_t stream, const AccumArgs& acc) {
    using L = InferSmem<ENC == 0 ? 80 : 64>;
    const uint32_t n_tiles = (n + kTile - 1) / kTile;
    const uint32_t grid = min((uint32_t)num_sms, (n_tiles + kGroups - 1) / kGroups);
    nrc_infer_kernel<ENC, ACC16><<<grid, kThreads, acc.out ? L::kTotal : L::kTotalNoAcc, stream>>>(in, out, n, params, w_img, gm, acc);
    return cudaGetLastError();
}

cudaError_t configure_infer_kernels() {   // per device (the attribute is not per process): called when a handle is created
    cudaError_t e = cudaFuncSetAttribute(nrc_infer_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, InferSmem<80>::kTotal);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(nrc_infer_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, InferSmem<80>::kTotal);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(nrc_infer_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, InferSmem<64>::kTotal);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(nrc_infer_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, InferSmem<64>::kTotal);
    return e;
}

cudaError_t launch_infer(int encoding, int acc16, const float* in, float* out, uint32_t n, const __half* params, const __half* w_img,
                         const GridMeta& gm, int num_sms, cudaStream_t stream, const AccumArgs* acc) {
    if (n == 0) return cudaSuccess;
    const AccumArgs a = acc ? *acc : AccumArgs{nullptr, nullptr, 0, 0, 0.0f};
    if (encoding == 0)
        return acc16 ? launch_infer_t<0, true>(in, out, n, params, w_img, gm, num_sms, stream, a)
                     : launch_infer_t<0, false>(in, out, n, params, w_img, gm, num_sms, stream, a);
    return acc16 ? launch_infer_t<1, true>(in, out, n, params, w_img, gm, num_sms, stream, a)
                 : launch_infer_t<1, false>(in, out, n, params, w_img, gm, num_sms, stream, a);
}

}  // namespace nrcb
