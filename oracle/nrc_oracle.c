/* nrc_oracle.c -- CPU ORACLE (test infrastructure, see nrc_oracle.h).  Each function cites the reference
 * file:line it restates ("tcnn/" = /root/reference/3rdparty/tiny-cuda-nn/).  Plain C, no intrinsics, no BLAS. */
#include "nrc_oracle.h"
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

typedef _Float16 h16;

uint16_t nrc_oracle_f2h(float f) { h16 h = (h16)f; uint16_t u; memcpy(&u, &h, 2); return u; }
/* fp16 -> fp32 is exact; a 65536-entry table keeps the scalar baseline from paying a soft-float call per weight. */
static float g_h2f[65536];
__attribute__((constructor)) static void init_h2f(void) {
    for (uint32_t u = 0; u < 65536; ++u) { uint16_t v = (uint16_t)u; h16 h; memcpy(&h, &v, 2); g_h2f[u] = (float)h; }
}
float nrc_oracle_h2f(uint16_t u) { return g_h2f[u]; }
/* MLP weights widened to fp32 once per call (exact), so the inner loops are plain float MACs. */
static float* widen_weights(const uint16_t* P, int n) {
    float* w = (float*)malloc((size_t)n * sizeof(float));
    for (int i = 0; i < n; ++i) w[i] = g_h2f[P[i]];
    return w;
}
static inline float rh(float f) { return (float)(h16)f; } /* round through fp16 */

/* ---------------------------------------------------------------- model layout */
/* grid.h:692-723 (offset table), common_device.h:856-868 (grid_scale / grid_resolution),
 * NRCNetworkConfigs.h:98-104 (base 16, per_level_scale 2, 16 levels, 2 features, log2 table 15). */
static float grid_scale(uint32_t level) { return exp2f((float)level * 1.0f) * 16.0f - 1.0f; }
static uint32_t grid_resolution(float scale) { return (uint32_t)ceilf(scale) + 1; }

int nrc_oracle_make_spec(int encoding, int log2_hashmap, nrc_oracle_spec* s) {
    memset(s, 0, sizeof(*s));
    if (encoding != NRC_ENC_FREQUENCY && encoding != NRC_ENC_HASH) return -1; /* NRCNetworkConfigs.h:129-131 */
    s->encoding = encoding;
    s->log2_hashmap = log2_hashmap > 0 ? log2_hashmap : 15;
    /* Freq: 36 + 24 + 6 = 66 -> 80; Hash: 32 + 24 + 6 = 62 -> 64 (network_with_input_encoding.h:46-47) */
    s->enc_width = encoding == NRC_ENC_FREQUENCY ? 80 : 64;
    s->n_mlp_params = 64 * s->enc_width + 4 * 64 * 64 + 16 * 64; /* fully_fused_mlp.cu:654-677 */
    if (encoding == NRC_ENC_HASH) {
        s->n_levels = 16;
        uint32_t off = 0;
        for (int l = 0; l < 16; ++l) {
            uint32_t res = grid_resolution(grid_scale((uint32_t)l));
            uint32_t max_params = 0xFFFFFFFFu / 2;
            uint32_t p = powf((float)res, 3) > (float)max_params ? max_params : res * res * res;
            p = (p + 7u) / 8u * 8u;
            uint32_t cap = 1u << s->log2_hashmap;
            if (p > cap) p = cap;
            s->level_offset[l] = off;
            off += p;
        }
        s->level_offset[16] = off;
        s->n_grid_params = (int)(off * 2);
    }
    s->n_params = s->n_mlp_params + s->n_grid_params;
    return 0;
}

/* ---------------------------------------------------------------- encodings */
/* tcnn/include/tiny-cuda-nn/encodings/triangle_wave.h:61-76 */
float nrc_oracle_triwave(float x_in, int log2_frequency) {
    const float x = scalbnf(x_in, log2_frequency - 1);
    const float val = x + (float)log2_frequency * 0.25f;
    return fabsf(val - floorf(val) - 0.5f) * 4 - 1;
}

/* tcnn/include/tiny-cuda-nn/common_device.h:915-920 */
float nrc_oracle_quartic_cdf(float x, float inv_radius) {
    const float u = x * inv_radius;
    const float u2 = u * u;
    const float u4 = u2 * u2;
    return fmaxf(0.0f, fminf(1.0f, ((float)15 / 16) * u * (1 - ((float)2 / 3) * u2 + ((float)1 / 5) * u4) + 0.5f));
}
#define QCDF nrc_oracle_quartic_cdf

/* AoS path used by the Frequency config: oneblob.h:47-67 (one_blob_subwarp_aligned, n_bins = 4).  Lane k
 * evaluates the wrapped CDF at the left boundary k/4 and takes the right one from lane k+1; the last bin
 * reads lane 0 (shuffle width 4 wraps) and adds 1. */
void nrc_oracle_oneblob4_aos(float x, float out[4]) {
    float left[4];
    for (int k = 0; k < 4; ++k) {
        const float lb = scalbnf((float)k, -2);
        left[k] = QCDF(lb - x, 4) + QCDF(lb - x - 1.0f, 4) + QCDF(lb - x + 1.0f, 4);
    }
    for (int k = 0; k < 4; ++k) {
        float right = left[(k + 1) & 3];
        if (k == 3) right += 1;
        out[k] = right - left[k];
    }
}
/* SoA path used by the Hash config (composite output layout follows the grid): oneblob.h:99-127. */
void nrc_oracle_oneblob4_soa(float x, float out[4]) {
    float left = QCDF(-x, 4) + QCDF(-x - 1.0f, 4) + QCDF(-x + 1.0f, 4);
    for (int k = 0; k < 4; ++k) {
        const float rb = scalbnf((float)(k + 1), -2);
        const float right = QCDF(rb - x, 4) + QCDF(rb - x - 1.0f, 4) + QCDF(rb - x + 1.0f, 4);
        out[k] = right - left;
        left = right;
    }
}

/* common_device.h:651-707 (coherent_prime_hash, grid_index with GridType::Hash) */
uint32_t nrc_oracle_grid_index(uint32_t hashmap_size, uint32_t resolution, uint32_t gx, uint32_t gy, uint32_t gz) {
    const uint32_t pg[3] = {gx, gy, gz};
    uint32_t stride = 1, index = 0;
    for (uint32_t dim = 0; dim < 3 && stride <= hashmap_size; ++dim) {
        index += pg[dim] * stride;
        stride *= resolution;
    }
    if (hashmap_size < stride) index = (gx * 1u) ^ (gy * 2654435761u) ^ (gz * 805459861u);
    return index % hashmap_size;
}

/* grid.h:94-169 (kernel_grid, Linear interpolation, half2 features, fma in half precision) */
static void grid_encode(const nrc_oracle_spec* s, const uint16_t* grid_h, const float* xyz, uint16_t* out32) {
    for (uint32_t level = 0; level < 16; ++level) {
        const uint16_t* grid = grid_h + (size_t)s->level_offset[level] * 2;
        const uint32_t hashmap_size = s->level_offset[level + 1] - s->level_offset[level];
        const float scale = grid_scale(level);
        const uint32_t resolution = grid_resolution(scale);
        float pos[3];
        uint32_t pg[3];
        for (int d = 0; d < 3; ++d) { /* pos_fract, common_device.h:856-868 */
            pos[d] = fmaf(scale, xyz[d], 0.5f);
            const float tmp = floorf(pos[d]);
            pg[d] = (uint32_t)(int)tmp;
            pos[d] -= tmp;
        }
        h16 r0 = 0, r1 = 0;
        for (uint32_t idx = 0; idx < 8; ++idx) {
            float w = 1;
            uint32_t pl[3];
            for (int d = 0; d < 3; ++d) {
                if ((idx & (1u << d)) == 0) { w *= 1 - pos[d]; pl[d] = pg[d]; }
                else                        { w *= pos[d];     pl[d] = pg[d] + 1; }
            }
            const uint32_t i = nrc_oracle_grid_index(hashmap_size, resolution, pl[0], pl[1], pl[2]) * 2;
            h16 wh = (h16)w, f0, f1;
            memcpy(&f0, &grid[i], 2);
            memcpy(&f1, &grid[i + 1], 2);
            r0 = (h16)((double)wh * (double)f0 + (double)r0); /* __hfma2: single rounding */
            r1 = (h16)((double)wh * (double)f1 + (double)r1);
        }
        memcpy(&out32[level * 2], &r0, 2);
        memcpy(&out32[level * 2 + 1], &r1, 2);
    }
}

/* composite.h:214-280: [pos encoding | one-blob(6 dims x 4) | identity(6) | 1.0 padding] (identity.h:59-67) */
static void encode_one(const nrc_oracle_spec* s, const uint16_t* params_h, const float* q, uint16_t* e) {
    int o = 0;
    if (s->encoding == NRC_ENC_FREQUENCY) {
        for (int d = 0; d < 3; ++d)
            for (int f = 0; f < 12; ++f) e[o++] = nrc_oracle_f2h(nrc_oracle_triwave(q[d], f));
    } else {
        grid_encode(s, params_h + s->n_mlp_params, q, e);
        o = 32;
    }
    for (int d = 3; d < 9; ++d) {
        float b[4];
        if (s->encoding == NRC_ENC_FREQUENCY) nrc_oracle_oneblob4_aos(q[d], b);
        else nrc_oracle_oneblob4_soa(q[d], b);
        for (int k = 0; k < 4; ++k) e[o++] = nrc_oracle_f2h(b[k]);
    }
    for (int d = 9; d < 15; ++d) e[o++] = nrc_oracle_f2h(q[d] * 1.0f + 0.0f);
    while (o < s->enc_width) e[o++] = nrc_oracle_f2h(1.0f);
}

void nrc_oracle_encode(const nrc_oracle_spec* s, const uint16_t* params_h, const float* in, uint32_t n,
                       uint16_t* enc_out) {
    for (uint32_t i = 0; i < n; ++i) encode_one(s, params_h, in + (size_t)i * 15, enc_out + (size_t)i * s->enc_width);
}

/* ---------------------------------------------------------------- MLP forward */
/* Semantics of kernel_mlp_fused (tcnn/src/fully_fused_mlp.cu:47-129,422-476,499-557): six bias-free matrices
 * W[out][in] row-major (fully_fused_mlp.cu:854-863), ReLU after every layer including the output
 * (NRCNetworkConfigs.h:26-33), every layer's activations stored as fp16. */
static void dense_relu(const float* W, int n_out, int n_in, const float* x, float* y, int acc_mode) {
    for (int o = 0; o < n_out; ++o) {
        const float* w = W + (size_t)o * n_in;
        float acc = 0;
        if (acc_mode == 0) {
            for (int k = 0; k < n_in; ++k) acc += w[k] * x[k];
        } else { /* accumulator kept in fp16 between 16-wide k blocks (wmma accumulator<__half>, :68) */
            for (int kb = 0; kb < n_in; kb += 16) {
                float blk = 0;
                for (int k = kb; k < kb + 16; ++k) blk += w[k] * x[k];
                acc = rh(acc + blk);
            }
        }
        y[o] = rh(acc > 0 ? acc : 0.0f);
    }
}

/* acts: [x(E) | h1..h5 (5x64) | y(16)] as floats holding fp16 values */
static void forward_one(const nrc_oracle_spec* s, const uint16_t* P, const float* Wf, const float* q, float* acts,
                        int acc_mode) {
    const int E = s->enc_width;
    uint16_t e[80];
    encode_one(s, P, q, e);
    for (int i = 0; i < E; ++i) acts[i] = nrc_oracle_h2f(e[i]);
    const float* W = Wf;
    dense_relu(W, 64, E, acts, acts + E, acc_mode);
    W += 64 * E;
    for (int l = 0; l < 4; ++l) {
        dense_relu(W, 64, 64, acts + E + 64 * l, acts + E + 64 * (l + 1), acc_mode);
        W += 64 * 64;
    }
    dense_relu(W, 16, 64, acts + E + 64 * 4, acts + E + 64 * 5, acc_mode);
}

typedef struct {
    const nrc_oracle_spec* s; const uint16_t* P; const float* Wf; const float* in; float* out; uint32_t lo, hi; int acc_mode;
} infer_job;
static void* infer_worker(void* a) {
    infer_job* j = (infer_job*)a;
    float acts[80 + 5 * 64 + 16];
    for (uint32_t i = j->lo; i < j->hi; ++i) {
        forward_one(j->s, j->P, j->Wf, j->in + (size_t)i * 15, acts, j->acc_mode);
        const float* y = acts + j->s->enc_width + 5 * 64;
        /* trim_and_cast: common_device.h:990-999 (first 3 of 16 lanes -> f32) */
        for (int c = 0; c < 3; ++c) j->out[(size_t)i * 3 + c] = y[c];
    }
    return 0;
}

/* nrc::Network::infer (NRCNetwork.cu:64-83) -> NetworkWithInputEncoding::inference (EMA weights are passed
 * by the caller as params_h).  Unlike the reference no rows past n are touched. */
void nrc_oracle_infer(const nrc_oracle_spec* s, const uint16_t* params_h, const float* in, uint32_t n,
                      float* out, int acc_mode, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    infer_job jobs[256];
    float* Wf = widen_weights(params_h, s->n_mlp_params);
    for (int t = 0; t < nthreads; ++t) {
        jobs[t] = (infer_job){s, params_h, Wf, in, out, (uint32_t)((uint64_t)n * t / nthreads),
                              (uint32_t)((uint64_t)n * (t + 1) / nthreads), acc_mode};
        if (nthreads == 1) infer_worker(&jobs[t]);
        else pthread_create(&th[t], 0, infer_worker, &jobs[t]);
    }
    if (nthreads > 1) for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    free(Wf);
}

/* ---------------------------------------------------------------- training */
/* 0: grid gradients summed in fp32 (default); 1: emulation of the reference's fp16 atomics, in sample order per thread. */
static int g_grid_grad_mode = 0;
void nrc_oracle_set_grid_grad_mode(int mode) { g_grid_grad_mode = mode; }
/* accumulator emulation of the TRAINING forward pass (0 fp32, 1 fp16 per 16-wide k block like the reference's wmma fragments) */
static int g_train_acc_mode = 0;
void nrc_oracle_set_train_acc_mode(int mode) { g_train_acc_mode = mode; }

typedef struct {
    nrc_oracle_state* st; const float* Wf; const float* in; const float* tgt; uint32_t lo, hi, n_norm;
    double* gw;   /* [n_mlp_params] partial weight gradients (loss-scaled)    */
    float* gg;    /* [n_grid_params] partial grid gradients (Hash) or NULL    */
    double loss;
} grad_job;

/* forward (trainer.h:97-141) + RelativeL2Luminance (relative_l2_luminance.h:54-86) + output-ReLU transfer
 * (common_device.h:241-249,747-757) + backward chain (fully_fused_mlp.cu:150-259) + weight gradients
 * (fully_fused_mlp.cu:785,819,828; dW = dY * X^T) + grid scatter (grid.h:252-255,301-319). */
static void* grad_worker(void* a) {
    grad_job* j = (grad_job*)a;
    const nrc_oracle_spec* s = &j->st->spec;
    const uint16_t* P = j->st->params;
    const int E = s->enc_width;
    const float loss_scale = j->st->loss_scale;
    const uint32_t n_total = j->n_norm * 3; /* n_elements / stride * dims */
    float acts[80 + 5 * 64 + 16], g[16], d[5][64], dx[64];
    for (uint32_t i = j->lo; i < j->hi; ++i) {
        const float* q = j->in + (size_t)i * 15;
        forward_one(s, P, j->Wf, q, acts, g_train_acc_mode);
        const float* x = acts;
        const float* h = acts + E; /* h[l*64 + k], l = 0..4 */
        const float* y = acts + E + 5 * 64;
        const float lum = 0.299f * y[0] + 0.587f * y[1] + 0.114f * y[2];
        const float den = lum * lum + 0.01f;
        for (int c = 0; c < 16; ++c) g[c] = 0;
        for (int c = 0; c < 3; ++c) {
            const float diff = y[c] - j->tgt[(size_t)i * 3 + c];
            j->loss += (double)(diff * diff / den / 1.0f / (float)n_total);
            const float gradient = 2 * diff / den / 1.0f;
            float gh = rh(loss_scale * gradient / (float)n_total);
            g[c] = rh(gh * (y[c] > 0 ? 1.0f : 0.0f));
        }
        /* delta_5 = (W_out^T g) .* [h5 > 0] */
        const float* Wout = j->Wf + 64 * E + 4 * 4096;
        for (int k = 0; k < 64; ++k) {
            float acc = 0;
            for (int c = 0; c < 16; ++c) acc += Wout[c * 64 + k] * g[c];
            d[4][k] = rh(h[4 * 64 + k] > 0 ? acc : 0.0f);
        }
        for (int l = 3; l >= 0; --l) { /* delta_l = (W_h(l+1)^T delta_(l+1)) .* [h_l > 0] */
            const float* W = j->Wf + 64 * E + l * 4096;
            for (int k = 0; k < 64; ++k) {
                float acc = 0;
                for (int o = 0; o < 64; ++o) acc += W[o * 64 + k] * d[l + 1][o];
                d[l][k] = rh(h[l * 64 + k] > 0 ? acc : 0.0f);
            }
        }
        /* weight gradients */
        double* gw = j->gw;
        for (int o = 0; o < 64; ++o) {
            const double dv = d[0][o];
            if (dv != 0) for (int k = 0; k < E; ++k) gw[o * E + k] += dv * x[k];
        }
        gw += 64 * E;
        for (int l = 0; l < 4; ++l, gw += 4096)
            for (int o = 0; o < 64; ++o) {
                const double dv = d[l + 1][o];
                if (dv != 0) for (int k = 0; k < 64; ++k) gw[o * 64 + k] += dv * h[l * 64 + k];
            }
        for (int c = 0; c < 3; ++c) {
            const double dv = g[c];
            if (dv != 0) for (int k = 0; k < 64; ++k) gw[c * 64 + k] += dv * h[4 * 64 + k];
        }
        if (j->gg) { /* Hash: dL/d(encoded) = W_in^T delta_1 (fp16), rows 0..31, then scatter to the 8 corners */
            for (int k = 0; k < 32; ++k) {
                float acc = 0;
                for (int o = 0; o < 64; ++o) acc += j->Wf[o * E + k] * d[0][o];
                dx[k] = rh(acc);
            }
            for (uint32_t level = 0; level < 16; ++level) {
                float* grid = j->gg + (size_t)s->level_offset[level] * 2;
                const uint32_t hashmap_size = s->level_offset[level + 1] - s->level_offset[level];
                const float scale = grid_scale(level);
                const uint32_t resolution = grid_resolution(scale);
                float pos[3]; uint32_t pg[3];
                for (int dd = 0; dd < 3; ++dd) {
                    pos[dd] = fmaf(scale, q[dd], 0.5f);
                    const float tmp = floorf(pos[dd]);
                    pg[dd] = (uint32_t)(int)tmp;
                    pos[dd] -= tmp;
                }
                for (uint32_t idx = 0; idx < 8; ++idx) {
                    float w = 1; uint32_t pl[3];
                    for (int dd = 0; dd < 3; ++dd) {
                        if ((idx & (1u << dd)) == 0) { w *= 1 - pos[dd]; pl[dd] = pg[dd]; }
                        else                         { w *= pos[dd];     pl[dd] = pg[dd] + 1; }
                    }
                    const uint32_t gi = nrc_oracle_grid_index(hashmap_size, resolution, pl[0], pl[1], pl[2]) * 2;
                    const float wh = rh(w);
                    if (g_grid_grad_mode == 0) {           /* fp32 accumulation of the exact products (what the CUDA kernels do) */
                        grid[gi] += wh * dx[level * 2];
                        grid[gi + 1] += wh * dx[level * 2 + 1];
                    } else {   /* the reference's arithmetic: (GRAD_T)weight * grad in half, atomicAdd(__half2) (grid.h:252-255) */
                        grid[gi] = rh(grid[gi] + rh(wh * dx[level * 2]));
                        grid[gi + 1] = rh(grid[gi + 1] + rh(wh * dx[level * 2 + 1]));
                    }
                }
            }
        }
    }
    return 0;
}

float nrc_oracle_grad_only(nrc_oracle_state* st, const float* in, const float* targets, uint32_t n, uint32_t n_norm,
                           float* grad_accum, int nthreads) {
    const nrc_oracle_spec* s = &st->spec;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 64) nthreads = 64;
    if ((uint32_t)nthreads > n) nthreads = n ? (int)n : 1;
    pthread_t th[64];
    grad_job jobs[64];
    float* Wf = widen_weights(st->params, s->n_mlp_params);
    for (int t = 0; t < nthreads; ++t) {
        jobs[t] = (grad_job){st, Wf, in, targets, (uint32_t)((uint64_t)n * t / nthreads),
                             (uint32_t)((uint64_t)n * (t + 1) / nthreads), n_norm,
                             (double*)calloc(s->n_mlp_params, sizeof(double)),
                             s->n_grid_params ? (float*)calloc(s->n_grid_params, sizeof(float)) : 0, 0.0};
        if (nthreads == 1) grad_worker(&jobs[t]);
        else pthread_create(&th[t], 0, grad_worker, &jobs[t]);
    }
    double loss = 0;
    if (nthreads > 1) for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    for (int i = 0; i < s->n_params; ++i) {
        double acc = 0;
        for (int t = 0; t < nthreads; ++t)
            acc = i < s->n_mlp_params ? acc + jobs[t].gw[i]
                  : g_grid_grad_mode ? (double)rh((float)acc + jobs[t].gg[i - s->n_mlp_params]) : acc + (double)jobs[t].gg[i - s->n_mlp_params];
        grad_accum[i] += (float)acc;
    }
    for (int t = 0; t < nthreads; ++t) { loss += jobs[t].loss; free(jobs[t].gw); free(jobs[t].gg); }
    free(Wf);
    return (float)loss;
}

/* Gradients become fp16 (the reference's gradient buffer is half, trainer.h:322-336), then
 * adam_step (adam.h:75-118) and ema_step_half_precision (ema.h:62-76,102-135). */
void nrc_oracle_apply_grads(nrc_oracle_state* st, const float* grad_accum) {
    const nrc_oracle_spec* s = &st->spec;
    st->step += 1;
    const float ema_debias_old = 1 - (float)pow((double)st->ema_decay, (double)(st->step - 1));
    const float ema_debias_new = 1.0f / (1 - (float)pow((double)st->ema_decay, (double)st->step));
    for (int i = 0; i < s->n_params; ++i) {
        st->grads[i] = nrc_oracle_f2h(grad_accum[i]);
        float gradient = nrc_oracle_h2f(st->grads[i]) / st->loss_scale;
        const int is_matrix = i < s->n_mlp_params;
        int skip = (!is_matrix && gradient == 0);
        if (!skip) {
            const float weight_fp = st->master[i];
            if (is_matrix) gradient += st->l2_reg * weight_fp;
            const float gradient_sq = gradient * gradient;
            const float m1 = st->m1[i] = st->beta1 * st->m1[i] + (1 - st->beta1) * gradient;
            const float m2 = st->m2[i] = st->beta2 * st->m2[i] + (1 - st->beta2) * gradient_sq;
            float lr = st->lr;
            const uint32_t cur = ++st->steps[i];
            lr *= sqrtf(1 - powf(st->beta2, (float)cur)) / (1 - powf(st->beta1, (float)cur));
            const float eff = fminf(fmaxf(lr / (sqrtf(m2) + st->eps), 0.0f), 3.402823466e+38f);
            const float new_weight = weight_fp - eff * m1;
            st->master[i] = new_weight;
            st->params[i] = nrc_oracle_f2h(new_weight);
        }
        const float filtered = (nrc_oracle_h2f(st->ema[i]) * st->ema_decay * ema_debias_old +
                                nrc_oracle_h2f(st->params[i]) * (1 - st->ema_decay)) * ema_debias_new;
        st->ema[i] = nrc_oracle_f2h(filtered);
    }
}

/* nrc::Network::train (NRCNetwork.cu:41-62) -> Trainer::training_step + Trainer::loss */
float nrc_oracle_train_step(nrc_oracle_state* st, const float* in, const float* targets, uint32_t n, int nthreads) {
    float* acc = (float*)calloc(st->spec.n_params, sizeof(float));
    const float loss = nrc_oracle_grad_only(st, in, targets, n, n, acc, nthreads);
    nrc_oracle_apply_grads(st, acc);
    free(acc);
    return loss;
}

/* ---------------------------------------------------------------- state */
nrc_oracle_state* nrc_oracle_state_create(int encoding, int log2_hashmap) {
    nrc_oracle_state* st = (nrc_oracle_state*)calloc(1, sizeof(*st));
    if (nrc_oracle_make_spec(encoding, log2_hashmap, &st->spec)) { free(st); return 0; }
    const size_t n = st->spec.n_params;
    st->master = (float*)calloc(n, 4); st->params = (uint16_t*)calloc(n, 2); st->ema = (uint16_t*)calloc(n, 2);
    st->grads = (uint16_t*)calloc(n, 2); st->m1 = (float*)calloc(n, 4); st->m2 = (float*)calloc(n, 4);
    st->steps = (uint32_t*)calloc(n, 4);
    /* NRCNetworkConfigs.h:18-24,43-47,92-97 + adam.h defaults */
    st->lr = encoding == NRC_ENC_FREQUENCY ? 1e-3f : 1e-2f;
    st->beta1 = 0.9f; st->beta2 = 0.999f;
    st->eps = encoding == NRC_ENC_FREQUENCY ? 1e-8f : 1e-15f;
    st->l2_reg = 1e-6f; st->ema_decay = 0.99f; st->loss_scale = 128.0f; /* common.h:232 */
    return st;
}
void nrc_oracle_state_destroy(nrc_oracle_state* st) {
    if (!st) return;
    free(st->master); free(st->params); free(st->ema); free(st->grads); free(st->m1); free(st->m2); free(st->steps);
    free(st);
}
void nrc_oracle_state_set_master(nrc_oracle_state* st, const float* master, int reset_optimizer) {
    const size_t n = st->spec.n_params;
    memcpy(st->master, master, n * 4);
    for (size_t i = 0; i < n; ++i) st->params[i] = nrc_oracle_f2h(master[i]); /* trainer.h:83-86 */
    if (reset_optimizer) {
        memset(st->ema, 0, n * 2); memset(st->m1, 0, n * 4); memset(st->m2, 0, n * 4); memset(st->steps, 0, n * 4);
        st->step = 0;
    }
}
static uint64_t splitmix64(uint64_t* s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static float u01(uint64_t* s) { return (float)(splitmix64(s) >> 40) * (1.0f / 16777216.0f); }
/* xavier uniform, scale = sqrt(6/(fan_in+fan_out)) (gpu_matrix.h:284-299); grid U(-1e-4, 1e-4) (grid.h:1059-1062).
 * The random stream is this oracle's own (splitmix64), not tcnn's pcg32. */
void nrc_oracle_state_init(nrc_oracle_state* st, uint64_t seed) {
    const nrc_oracle_spec* s = &st->spec;
    float* m = (float*)malloc((size_t)s->n_params * 4);
    uint64_t rng = seed;
    const int E = s->enc_width;
    const int rows[6] = {64, 64, 64, 64, 64, 16}, cols[6] = {E, 64, 64, 64, 64, 64};
    size_t o = 0;
    for (int l = 0; l < 6; ++l) {
        const float scale = sqrtf(6.0f / (float)(rows[l] + cols[l]));
        for (int i = 0; i < rows[l] * cols[l]; ++i) m[o++] = u01(&rng) * 2 * scale - scale;
    }
    for (; o < (size_t)s->n_params; ++o) m[o] = u01(&rng) * 2e-4f - 1e-4f;
    nrc_oracle_state_set_master(st, m, 1);
    free(m);
}

/* ---------------------------------------------------------------- NEXT rows (frame helpers) */
/* nrc/shaders/nrc_helpers.cu:77-129 */
void nrc_oracle_accumulate(const float* radiance, const float* throughput, float* out4, uint32_t n, int mode,
                           uint32_t iteration_index) {
    const float acc_weight = 1.0f / (float)(iteration_index + 1);
    for (uint32_t i = 0; i < n; ++i) {
        const float* r = radiance + 3 * (size_t)i; const float* t = throughput + 3 * (size_t)i; float* o = out4 + 4 * (size_t)i;
        switch (mode) {
            case 0: for (int c = 0; c < 3; ++c) o[c] = fmaf(t[c] * r[c], acc_weight, o[c]); o[3] = 1.0f; break;   /* Full */
            case 2: for (int c = 0; c < 3; ++c) o[c] = r[c] * t[c]; o[3] = 1.0f; break;                          /* CacheOnly */
            case 4: for (int c = 0; c < 3; ++c) o[c] = r[c]; o[3] = 1.0f; break;
            case 5: for (int c = 0; c < 3; ++c) o[c] = t[c]; o[3] = 1.0f; break;
            default: break;                                                                                       /* NoCache, CacheFirstVertex */
        }
    }
}
/* nrc/shaders/nrc_helpers.cu:131-224 (USE_REFLECTANCE_FACTORING = 0, config.h:118) */
void nrc_oracle_propagate(const void* end_vertices, const float* end_radiance, const void* records, float* targets,
                          uint32_t n_tiles) {
    const int32_t* ev = (const int32_t*)end_vertices;
    const int32_t* rec = (const int32_t*)records;
    for (uint32_t tile = 0; tile < n_tiles; ++tile) {
        float mask; memcpy(&mask, &ev[4 * tile + 1], 4);
        float last[3];
        for (int c = 0; c < 3; ++c) last[c] = end_radiance[3 * (size_t)tile + c] * mask;
        int i = ev[4 * tile];
        while (i >= 0) {
            const int32_t* r = rec + 7 * (size_t)i;
            for (int c = 0; c < 3; ++c) {
                float lt; memcpy(&lt, &r[1 + c], 4);
                last[c] = fmaf(lt, last[c], targets[3 * (size_t)i + c]);
                targets[3 * (size_t)i + c] = last[c];
            }
            i = r[0];
        }
    }
}
static uint32_t feistel_round(uint32_t x, uint32_t key) {
    x = (x + key) * 0x9E3779B1u; x ^= x >> 15; x *= 0x85EBCA77u; x ^= x >> 13;
    return x & 0xFFu;
}
uint32_t nrc_oracle_permutation(uint32_t i, uint32_t seed) {
    uint32_t l = (i >> 8) & 0xFFu, r = i & 0xFFu;
    for (uint32_t k = 0; k < 4; ++k) {
        const uint32_t t = l ^ feistel_round(r, seed * 4u + k + 0x632BE5ABu * (k + 1u));
        l = r; r = t;
    }
    return (l << 8) | r;
}
/* nrc/shaders/nrc_helpers.cu:226-249 */
void nrc_oracle_permute(const float* q_src, const float* t_src, float* q_dst, float* t_dst, int num_records, uint32_t seed) {
    const int n = num_records < 65536 ? num_records : 65536;
    if (n <= 0) return;
    for (uint32_t d = 0; d < 65536; ++d) {
        const uint32_t s = nrc_oracle_permutation(d, seed) % (uint32_t)n;
        memcpy(q_dst + 15 * (size_t)d, q_src + 15 * (size_t)s, 60);
        memcpy(t_dst + 3 * (size_t)d, t_src + 3 * (size_t)s, 12);
    }
}
/* nrc/src/Device.cpp:818-828 */
void nrc_oracle_adjust_tile_size(int tx, int ty, int num_records, int* nx, int* ny) {
    const float ratio = (float)(num_records * 1.25f) / (float)65536;
    const float r = sqrtf(ratio);
    int x = (int)(tx * r + 0.5f), y = (int)(ty * r + 0.5f);
    *nx = x > 2 ? x : 2; *ny = y > 2 ? y : 2;
}
