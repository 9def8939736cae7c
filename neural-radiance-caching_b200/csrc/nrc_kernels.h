// Internal launch interface between the host side (nrc_b200.cu) and the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace nrcb {

struct GridMeta;

// nrc_infer.cu -- fused query kernel (encodings + 6-matrix MLP), EMA weights.
cudaError_t launch_infer(int encoding, const float* in, float* out, uint32_t n, const __half* params, const GridMeta& gm,
                         int num_sms, cudaStream_t stream);

// nrc_train.cu -- fused forward + loss + backward + per-tile weight-gradient kernel, and the
// gradient reduction / Adam / EMA kernels.
struct TrainBuffers {
    const __half* params;    // fp16 training weights
    float* partials;         // [n_tiles][n_mlp_params] per-tile weight-gradient partials (loss-scaled)
    float* grid_grad;        // [n_grid_params] fp32 grid gradient accumulator (Hash), zeroed by the optimizer kernel
    float* loss_partials;    // [n_tiles]
};
cudaError_t launch_train_fwd_bwd(int encoding, const float* in, const float* targets, uint32_t n_local, uint32_t n_global,
                                 const TrainBuffers& tb, const GridMeta& gm, cudaStream_t stream);

struct OptimState {
    float* master; __half* params; __half* ema; __half* grads; float* m1; float* m2; uint32_t* steps;
    uint32_t n_params, n_mlp;
    float lr, beta1, beta2, eps, l2_reg, ema_decay, loss_scale;
};
// Sums the per-tile partials (+ grid accumulator) into grad_f32[n_params]; writes loss_out = sum(loss_partials).
cudaError_t launch_reduce_grads(const TrainBuffers& tb, uint32_t n_tiles, uint32_t n_params, uint32_t n_mlp,
                                float* grad_f32, float* loss_out, cudaStream_t stream);
// grad_f32 -> fp16 gradient, Adam (adam.h:75-118), EMA (ema.h:62-76); clears grid_grad for the next step.
cudaError_t launch_adam_ema(const OptimState& os, const float* grad_f32, float* grid_grad_to_clear, uint32_t step,
                            cudaStream_t stream);

}  // namespace nrcb
