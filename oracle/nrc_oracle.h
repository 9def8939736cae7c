/* nrc_oracle.h -- CPU ORACLE for the NRC cache path.  TEST INFRASTRUCTURE ONLY.
 *
 * A scalar C restatement of the reference's algorithm (tiny-cuda-nn kernels reached from
 * nrc::Network, /root/reference/nrc/src/NRCNetwork.cu:41-83).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may link or call this; the product
 * (neural-radiance-caching_b200/) never does.
 *
 * PARITY PIN: the reference ships no golden vectors / known-answer tests for this path
 * (SURVEY.md section 4), so this oracle is pinned by (a) closed-form checks derived from the reference
 * source (tests/test_oracle.py) and (b) outputs of the reference's own tcnn build run on a B200
 * (oracle/_ref, fixtures under tests/golden/ with the generating script).  Until (b) is committed
 * the status is "parity unpinned" -- see DESIGN.md.
 */
#ifndef NRC_ORACLE_H
#define NRC_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { NRC_ENC_FREQUENCY = 0, NRC_ENC_HASH = 1 };   /* nrc::InputEncoding, neural_radiance_caching.h:24-27 */
enum { NRC_IN_DIMS = 15, NRC_OUT_DIMS = 3, NRC_WIDTH = 64, NRC_OUT_PAD = 16, NRC_BATCH = 16384 };

typedef struct nrc_oracle_spec {
    int encoding;
    int enc_width;          /* E: 80 (Frequency) / 64 (Hash)                                   */
    int n_mlp_params;       /* 64*E + 4*64*64 + 16*64                                          */
    int n_grid_params;      /* 0 / sum(level sizes)*2                                          */
    int n_params;
    int n_levels;           /* 16                                                              */
    int log2_hashmap;       /* 15 by default                                                   */
    uint32_t level_offset[17]; /* in grid *entries* (2 features each), cumulative               */
} nrc_oracle_spec;

/* Layout facts of the reference model (NRCNetworkConfigs.h:37-135, grid.h:692-723). */
int nrc_oracle_make_spec(int encoding, int log2_hashmap, nrc_oracle_spec* out);

/* Encoded features exactly as CompositeEncoding::forward_impl stores them (fp16 bits), sample-major [n][E]. */
void nrc_oracle_encode(const nrc_oracle_spec* s, const uint16_t* params_h, const float* in, uint32_t n,
                       uint16_t* enc_out);

/* Full query: encodings + 6-matrix ReLU MLP, out = float[n][3].  acc_mode 0: fp32 accumulate, activations
 * rounded to fp16 per layer (what the B200 kernels do); 1: additionally round the accumulator to fp16 after
 * every 16-wide k block (approximates the reference's wmma accumulator<__half>).  nthreads >= 1. */
void nrc_oracle_infer(const nrc_oracle_spec* s, const uint16_t* params_h, const float* in, uint32_t n,
                      float* out, int acc_mode, int nthreads);

/* Trainer state: every array has n_params entries. */
typedef struct nrc_oracle_state {
    nrc_oracle_spec spec;
    float* master;      /* fp32 weights (Trainer::m_params_full_precision)       */
    uint16_t* params;   /* fp16 training weights                                  */
    uint16_t* ema;      /* fp16 EMA weights = inference weights (ema.h:93-156)    */
    uint16_t* grads;    /* fp16 gradients of the last step (loss-scaled by 128)   */
    float* m1;          /* Adam first moments                                     */
    float* m2;          /* Adam second moments                                    */
    uint32_t* steps;    /* per-parameter step counters (adam.h:106)               */
    uint32_t step;      /* global step                                            */
    float lr, beta1, beta2, eps, l2_reg, ema_decay, loss_scale;
} nrc_oracle_state;

nrc_oracle_state* nrc_oracle_state_create(int encoding, int log2_hashmap);
void nrc_oracle_state_destroy(nrc_oracle_state* st);
/* Deterministic xavier-uniform init (own RNG; parity tests copy weights between implementations). */
void nrc_oracle_state_init(nrc_oracle_state* st, uint64_t seed);
/* Replace weights (master fp32; fp16 copy derived by rounding), keep/zero optimizer state. */
void nrc_oracle_state_set_master(nrc_oracle_state* st, const float* master, int reset_optimizer);

/* One optimizer step on n samples (the reference uses n = 16384): forward with the raw fp16 weights,
 * RelativeL2Luminance loss, backward, weight gradients, Adam + EMA.  Returns the loss (sum of per-element
 * values, trainer.h:205-207).  The gradient/loss normalisation uses n_total = 3*n_norm, with n_norm = n
 * unless the caller shards a batch (data-parallel emulation): then n_norm = global batch size and
 * `apply` = 0 only accumulates the fp32 gradient into grad_accum (n_params floats) without stepping. */
/* Hash grid-gradient accumulation: 0 = fp32 sums of the exact products (default), 1 = the reference's arithmetic:
 * (GRAD_T)weight * grad rounded to half, added with atomicAdd(__half2) (grid.h:252-255,301-319), emulated in sample order. */
void nrc_oracle_set_grid_grad_mode(int mode);
/* Accumulators of the training forward pass: 0 = fp32 (default), 1 = fp16 per 16-wide k block (fully_fused_mlp.cu:68). */
void nrc_oracle_set_train_acc_mode(int mode);
float nrc_oracle_train_step(nrc_oracle_state* st, const float* in, const float* targets, uint32_t n, int nthreads);
float nrc_oracle_grad_only(nrc_oracle_state* st, const float* in, const float* targets, uint32_t n, uint32_t n_norm,
                           float* grad_accum, int nthreads);
void nrc_oracle_apply_grads(nrc_oracle_state* st, const float* grad_accum);

/* ---- NEXT rows: CPU restatements of the helper kernels around the network.
 * PARITY: these reference kernels live in the OptiX module (nrc/shaders/nrc_helpers.cu reads launch state from the
 * `sysData` constant and needs optix.h / system_data.h), so they cannot be compiled without the OptiX SDK:
 * "parity unpinned" by reference outputs for these rows; they are restated from the source.  Multiply-adds are written
 * as fmaf because nvcc (default -fmad=true) contracts `a += b * c` in the reference kernels. */
/* accumulate_render_radiance, nrc_helpers.cu:77-129; out4 = float4[n] */
void nrc_oracle_accumulate(const float* radiance, const float* throughput, float* out4, uint32_t n, int mode,
                           uint32_t iteration_index);
/* propagate_train_radiance, nrc_helpers.cu:131-224; end_vertices = {int start; float mask; int; int}[n_tiles],
 * records = {int propTo; float lt[3]; int; int; int}[<=65536] */
void nrc_oracle_propagate(const void* end_vertices, const float* end_radiance, const void* records, float* targets,
                          uint32_t n_tiles);
/* The keyed 16-bit bijection that replaces cuRAND keys + CUB sort (NRCUtil.cu:20-35); must equal nrc_b200_permutation. */
uint32_t nrc_oracle_permutation(uint32_t i, uint32_t seed);
/* permute_train_data, nrc_helpers.cu:226-249 */
void nrc_oracle_permute(const float* q_src, const float* t_src, float* q_dst, float* t_dst, int num_records, uint32_t seed);
/* Device::adjustTileSize, Device.cpp:818-828 */
void nrc_oracle_adjust_tile_size(int tx, int ty, int num_records, int* nx, int* ny);

/* Pieces exposed for unit tests. */
float nrc_oracle_quartic_cdf(float x, float inv_radius);
void nrc_oracle_oneblob4_aos(float x, float out[4]);
void nrc_oracle_oneblob4_soa(float x, float out[4]);
float nrc_oracle_triwave(float x, int log2_frequency);
uint32_t nrc_oracle_grid_index(uint32_t hashmap_size, uint32_t resolution, uint32_t gx, uint32_t gy, uint32_t gz);
uint16_t nrc_oracle_f2h(float f);
float nrc_oracle_h2f(uint16_t h);

#ifdef __cplusplus
}
#endif
#endif
