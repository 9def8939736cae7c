// Internal launch interface between the host side (nrc_b200.cu) and the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace nrcb {

struct GridMeta;

// nrc_infer.cu -- fused query kernel (encodings + 6-matrix MLP), EMA weights.  Optional fused epilogue: the first
// acc->n queries are pixels whose radiance is accumulated into the output image (accumulate_render_radiance).
struct AccumArgs {
    const float* throughput;   // float3[n]
    float4* out;               // float4[n], nullptr = no accumulation
    uint32_t n;
    int mode;                  // nrc::RenderMode
    float weight;              // 1 / (iterationIndex + 1)
};
// params: fp16 parameter vector (the grid part is read from it); w_img: UMMA shared-memory image of its MLP part
// (brought in with one bulk copy per CTA).  acc16 = 1: fp16 accumulators (the arithmetic of the reference's
// wmma::accumulator<__half> fragments, fully_fused_mlp.cu:68) instead of fp32.
cudaError_t launch_infer(int encoding, int acc16, const float* in, float* out, uint32_t n, const __half* params, const __half* w_img,
                         const GridMeta& gm, int num_sms, cudaStream_t stream, const AccumArgs* acc = nullptr);
// cudaFuncAttributeMaxDynamicSharedMemorySize is per device: called for the current device when a handle is created.
cudaError_t configure_infer_kernels();
cudaError_t configure_train_kernels();

// nrc_train.cu -- ONE cooperative kernel per optimizer step:
//   phase 1  per 128-sample tile: encodings, forward, loss, backward, per-tile weight gradients (tcgen05)
//   phase 2  grid barrier
//   phase 3  every CTA reduces its slice of the gradient over all tiles (fixed order) and either applies
//            Adam + EMA to that slice (apply = 1) or leaves the fp32 gradient for a data-parallel all-reduce (apply = 0).
// The reference's per-frame loop of NUM_BATCHES train() calls (Device::nrcTrainRadiance) can run as one launch (n_steps).
struct TrainArgs {
    const float* in; const float* targets;
    uint32_t n_local, n_global;      // samples of this launch / of the global batch (loss normalisation)
    // parameter block
    float* master; __half* params; __half* params_img; __half* ema; __half* ema_img; __half* grads;
    float* m1; float* m2; uint32_t* steps;
    // scratch
    float* partials;                 // [n_tiles][n_mlp] per-tile weight-gradient partials (loss-scaled)
    float* grid_grad;                // [n_grid] fp32 scatter accumulator (Hash), cleared in phase 3
    float* loss_partials;            // [n_tiles]
    float* grad_f32;                 // [n_params] reduced fp32 gradient (apply = 0: the all-reduce payload)
    float* loss_out;
    uint32_t* barrier_counter;       // monotonically increasing arrival counters: [0] top (polled), [32 * (1 + g)] group g
    uint32_t barrier_target;         // value the top counter reaches when every CTA has arrived at the launch's FIRST barrier
    uint32_t n_params, n_mlp;
    float lr, beta1, beta2, eps, l2_reg, ema_decay, loss_scale;
    float inv_loss_scale;            // 1 / loss_scale, used instead of the division when loss_scale is a power of two (exact)
    int loss_scale_pow2;
    float ema_debias_old[4], ema_debias_new[4];   // per step of the launch
    int apply;
    int acc16;                       // forward / backward-chain MMAs with fp16 accumulators (the reference's wmma accumulator<__half>)
    int grid_grad_fp16;              // Hash: grid gradients as half products summed with half2 atomics (the reference's arithmetic)
    // n_steps optimizer steps in ONE launch (apply = 1): step s trains on samples [s*n_local, (s+1)*n_local)
    // of `in` / `targets` (fused shuffle: destination indices perm_base + s*n_local + ...), writes its loss to
    // loss_out[s]; two grid barriers per step (gradients complete / weights updated), the last one omitted.
    uint32_t n_steps;
    int wide_grid;                   // Hash on a 148-SM device: the launch has 148 CTAs (train_grid)
    long long* timeline;             // optional [n_tiles][8] clock64 stamps of the phases (debug), or nullptr
    // Data-parallel exchange fused into phase 3 (dp_world > 1, apply = 1): every rank pushes its slice gradients into
    // row `dp_rank` of EVERY peer's exchange buffer through NVLink peer pointers and sums the rows it received in rank
    // order (bit-identical replicas, no NCCL call).  MLP slice + loss: (value, step) pairs in 8-byte stores, polled by
    // the receiver; grid part: floats + fence + per-slice flag.
    //   xbuf[r]  : rank r's buffer  float[2 parities][dp_world][x_stride]   (xbuf[dp_rank] is local memory)
    //   xflag[r] : rank r's flags   u32  [2 parities][dp_world][128 slices] (value = step number)
    // Fused shuffle (perm_n > 0): sample d = perm_base + tile*128 + row reads record permute16(d, perm_seed) % perm_n
    // of `in` / `targets` (the reference's permute_train_data, nrc_helpers.cu:226-249, without the 4.7 MB copy).
    uint32_t perm_n, perm_seed, perm_base;
    uint32_t opt_step;               // number of the (first) optimizer step of this launch: 1, 2, ... (Adam/EMA bias corrections)
    uint32_t dp_world, dp_rank, x_stride;
    uint32_t dp_seq;                 // exchange sequence number of the (first) step: tag / flag value and buffer parity; counts
                                     // from 1 since the buffers were connected and is never reset (weights may be re-uploaded)
    float* xbuf[8];
    uint32_t* xflag[8];
    uint32_t* error_flag;            // host-mapped word: a peer that stays silent for dp_timeout_ms raises it (no trap: the
    uint32_t dp_timeout_ms;          // context survives, the next API call reports the error)
};
constexpr int kMaxDpRanks = 8;
#ifndef NRCB_BARRIER_TREE
#define NRCB_BARRIER_TREE 0      // measured: the two dependent L2 round trips cost more than the serialised arrivals (+1.3 us per barrier, profiles/r02_ab_variants.json)
#endif
// CTAs of a step launch: one per 128-sample tile, but never fewer than 128 -- short data-parallel shards keep the gradient
// reduction / optimizer phase 128 slices wide (the extra CTAs skip phase 1).  The grid barrier counts this many arrivals.
// Hash (wide = 1): 148 CTAs -- the 20 CTAs without a data tile only take part in phase 3, whose pass over the 991 232 grid
// parameters (26 B each) then runs on every SM of a B200 instead of 128 of them.
#ifndef NRCB_HASH_GRID148
#define NRCB_HASH_GRID148 1
#endif
constexpr uint32_t kWideGrid = 148;
constexpr uint32_t kDpFlagSlots = 160;   // per (parity, source rank): one flag per gradient slice (= CTA), >= the widest grid
inline __host__ __device__ uint32_t train_grid(uint32_t n_local, int wide = 0) {
    const uint32_t t = n_local / 128, g = t < 128 ? 128 : t;
    return (NRCB_HASH_GRID148 && !NRCB_BARRIER_TREE && wide && g < kWideGrid) ? kWideGrid : g;
}
// Two-level grid barrier (nrc_train.cu grid_arrive): CTAs arrive in groups of kBarrierFan on per-group counters, the last
// arriver of a group on the top counter; one barrier instance adds this many arrivals to the top counter.
constexpr uint32_t kBarrierFan = 16;
constexpr uint32_t kBarrierWords = 32 * (1 + 16);   // top counter + up to 16 group counters, 128 bytes apart
inline __host__ __device__ uint32_t train_barrier_arrivals(uint32_t n_local, int wide = 0) {
    return NRCB_BARRIER_TREE ? train_grid(n_local, wide) / kBarrierFan : train_grid(n_local, wide);
}
cudaError_t launch_train_step(int encoding, const TrainArgs& a, const GridMeta& gm, cudaStream_t stream);

// Adam + EMA from an already reduced fp32 gradient (after a data-parallel all-reduce); clears grid_grad.
cudaError_t launch_adam_ema(const TrainArgs& a, cudaStream_t stream);
// Rebuilds the UMMA shared-memory image of the fp16 training weights (after a host upload).
cudaError_t launch_make_image(const __half* params, __half* params_img, uint32_t enc_width, uint32_t n_mlp,
                              cudaStream_t stream);

// nrc_frame.cu -- the helper kernels either side of the network in the per-frame sequence
cudaError_t launch_accumulate(const float* radiance, const float* throughput, float* out4, uint32_t n_pixels, int mode,
                              uint32_t iteration_index, cudaStream_t stream);
cudaError_t launch_propagate(const void* end_vertices, const float* end_radiance, const void* records, float* targets,
                             uint32_t n_tiles, cudaStream_t stream);
cudaError_t launch_permute(const float* q_src, const float* t_src, float* q_dst, float* t_dst, uint32_t num_records,
                           uint32_t seed, cudaStream_t stream);
cudaError_t launch_permute_with(const float* q_src, const float* t_src, float* q_dst, float* t_dst, const int* permutation,
                                uint32_t num_records, cudaStream_t stream);

// measurement aid: read bandwidth of an L2-resident buffer (GB/s)
cudaError_t l2_read_probe(size_t bytes, int reps, float* gbs);

}  // namespace nrcb
