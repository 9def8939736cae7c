// Host side of the B200 neural radiance cache: implements the C ABI of include/nrc_b200.h.
// Mirrors nrc::Network (/root/reference/nrc/src/NRCNetwork.cu) + the parts of tcnn::Trainer it uses
// (tcnn/include/tiny-cuda-nn/trainer.h:50-95,163-207,322-336): one parameter block
// [fp32 master | fp16 weights | fp16 EMA | fp16 grads | Adam m, v, steps], everything preallocated at
// create (no arena, no per-call malloc, no per-step graph capture, no host sync unless loss_h != NULL).
#include "../../include/nrc_b200.h"
#include "nrc_kernels.h"
#include "nrc_encode.cuh"
#include "nrc_frame.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

using namespace nrcb;

static thread_local std::string g_last_error;
const char* nrc_b200_last_error(void) { return g_last_error.c_str(); }

static int fail(int code, const std::string& msg) { g_last_error = msg; return code; }
#define CU(x)                                                                                     \
    do {                                                                                          \
        cudaError_t e_ = (x);                                                                     \
        if (e_ != cudaSuccess)                                                                    \
            return fail(NRC_B200_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_));      \
    } while (0)

// One handle = one device + one stream (SURVEY.md 8b "Threading"): every entry point makes the handle's device current
// for the duration of the call and restores the caller's, so several handles on different GPUs can be driven from one
// thread (or one each from several threads).
struct DeviceGuard {
    int prev = -1; bool switched = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

struct nrc_b200_network {
    int encoding = 0, log2_hashmap = 15, device = 0, num_sms = 148;
    uint64_t seed = 1337;
    cudaStream_t stream = nullptr;
    uint32_t n_params = 0, n_mlp = 0, enc_width = 80;
    GridMeta gm{};
    // parameter block (device)
    float* master = nullptr; __half* params = nullptr; __half* ema = nullptr; __half* grads = nullptr;
    __half* params_img = nullptr;   // UMMA shared-memory image of the fp16 training MLP weights (bulk-copied by the step kernel)
    __half* ema_img = nullptr;      // the same image of the fp16 EMA (= inference) MLP weights (bulk-copied by the query kernel)
    uint32_t* barrier_counter = nullptr; uint32_t barrier_total = 0;
    long long* timeline = nullptr;   // debug: per-tile phase stamps of the step kernel
    // data-parallel exchange over NVLink peer memory (nrc_b200_dp_*): one IPC allocation [xbuf | xflag] per rank
    void* dp_local = nullptr; void* dp_peer[kMaxDpRanks] = {}; int dp_rank = 0, dp_world = 1;
    bool dp_shared_process = false;  // peers are handles of this process (peer access), not CUDA IPC mappings
    uint32_t x_stride = 0; size_t xbuf_bytes = 0;   // grid barrier of the cooperative step kernel
    float* m1 = nullptr; float* m2 = nullptr; uint32_t* steps = nullptr;
    // training scratch (device)
    float* partials = nullptr; float* grid_grad = nullptr; float* loss_partials = nullptr;
    float* grad_f32 = nullptr; float* loss_d = nullptr;
    float* loss_pinned = nullptr;   // 4 floats (one per batch of a frame)
    float* loss4_d = nullptr;
    // hyper-parameters (NRCNetworkConfigs.h:18-24,43-47,92-97; adam.h defaults)
    float lr = 1e-3f, beta1 = 0.9f, beta2 = 0.999f, eps = 1e-8f, l2_reg = 1e-6f, ema_decay = 0.99f, loss_scale = 128.0f;
    uint32_t step = 0;
    uint32_t pending_tiles = 0;  // tiles of the gradient waiting for train_apply
    // host-buffer inference pipeline
    static constexpr int kHostChunks = 3;
    float* pin_in[kHostChunks] = {}; float* pin_out[kHostChunks] = {};
    float* dev_in[kHostChunks] = {}; float* dev_out[kHostChunks] = {};
    cudaStream_t hstream[kHostChunks] = {}; cudaEvent_t hevent[kHostChunks] = {};
    uint32_t host_chunk = 0;
    float* train_in_d = nullptr; float* train_tgt_d = nullptr;   // staging for train_host (one 16384 batch)
    uint64_t launches = 0;
    bool destroyed = false;
    int acc16 = 0;                   // 1: query kernel with fp16 accumulators (the reference's wmma accumulator<__half> arithmetic)
    uint32_t dp_seq = 0;             // data-parallel exchange sequence number (tags / flags / buffer parity); never reset while connected
    uint32_t* err_host = nullptr;    // pinned, device-mapped word the kernels raise on a peer time-out (instead of a trap)
    uint32_t* err_dev = nullptr;
    uint32_t dp_timeout_ms = 30000;  // NRC_B200_DP_TIMEOUT_MS
};
typedef nrc_b200_network Net;

static int device_error(Net* h) {
    const uint32_t code = *(volatile uint32_t*)h->err_host;
    return fail(NRC_B200_ERR_CUDA, "data-parallel exchange: a peer did not deliver exchange step " + std::to_string(code) +
                                       " within " + std::to_string(h->dp_timeout_ms) + " ms (NRC_B200_DP_TIMEOUT_MS); the replicas are out of sync");
}
// Common prologue of the per-handle entry points: null -> invalid argument; destroyed -> the reference's silent no-op
// (NRCNetwork.cu:43,66), reported as NRC_B200_ERR_DESTROYED; the handle's device made current for the call; an error
// a kernel raised earlier (peer time-out) is reported instead of launching more work.
#define ENTER(h)                                                                   \
    if (!(h)) return fail(NRC_B200_ERR_INVALID, "null handle");                    \
    if ((h)->destroyed) return NRC_B200_ERR_DESTROYED;                             \
    DeviceGuard guard_((h)->device);                                               \
    if ((h)->err_host && *(volatile uint32_t*)(h)->err_host) return device_error(h)

static constexpr uint32_t kMaxTrainTiles = NRC_B200_BATCH_SIZE / 128;

// ---------------------------------------------------------------- model layout
static void make_layout(Net* n) {
    n->enc_width = n->encoding == NRC_B200_FREQUENCY ? 80 : 64;
    n->n_mlp = 64 * n->enc_width + 4 * 64 * 64 + 16 * 64;  // fully_fused_mlp.cu:654-677
    uint32_t off = 0;
    memset(&n->gm, 0, sizeof(n->gm));
    if (n->encoding == NRC_B200_HASH) {
        for (int l = 0; l < 16; ++l) {  // grid.h:692-723
            const float scale = exp2f((float)l) * 16.0f - 1.0f;  // common_device.h:856-861
            const uint32_t res = (uint32_t)ceilf(scale) + 1;
            const double dense = (double)res * res * res;   // params_in_level before the cap (grid.h:703-707)
            const uint32_t cap = 1u << n->log2_hashmap;
            uint32_t size = dense > (double)cap ? cap : (uint32_t)dense;
            size = (size + 7u) / 8u * 8u;
            if (size > cap) size = cap;
            n->gm.offset[l] = off;
            n->gm.scale[l] = scale;
            // grid_index (common_device.h:691-707), literally, in uint32: which dimensions enter the linear index with
            // which stride, and whether the level ends up hashed.  res^2 wraps to 0 for res >= 2^16, so fine levels of a
            // table with >= res entries leave the loop with stride 0 <= size: NOT hashed, index = x + y*res.
            uint32_t stride = 1, st[3] = {0, 0, 0};
            for (int dim = 0; dim < 3 && stride <= size; ++dim) { st[dim] = stride; stride *= res; }
            n->gm.stride_y[l] = st[1];
            n->gm.stride_z[l] = st[2];
            if (!(size < stride)) n->gm.dense_mask |= 1u << l;
            off += size;
        }
        n->gm.offset[16] = off;
    }
    n->n_params = n->n_mlp + off * 2;
}

static uint64_t splitmix64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static float u01(uint64_t& s) { return (float)(splitmix64(s) >> 40) * (1.0f / 16777216.0f); }

static void dp_teardown(Net* n) {
    for (int r = 0; r < kMaxDpRanks; ++r) {
        if (n->dp_peer[r] && r != n->dp_rank && !n->dp_shared_process) cudaIpcCloseMemHandle(n->dp_peer[r]);
        n->dp_peer[r] = nullptr;
    }
    cudaFree(n->dp_local);
    n->dp_local = nullptr; n->dp_world = 1; n->dp_rank = 0; n->dp_shared_process = false;
}

static void free_all(Net* n) {
    cudaFree(n->master); cudaFree(n->params); cudaFree(n->ema); cudaFree(n->grads);
    cudaFree(n->m1); cudaFree(n->m2); cudaFree(n->steps); cudaFree(n->params_img); cudaFree(n->ema_img); cudaFree(n->barrier_counter);
    n->params_img = n->ema_img = nullptr; n->barrier_counter = nullptr; n->barrier_total = 0;
    cudaFree(n->timeline); n->timeline = nullptr;
    dp_teardown(n);
    cudaFree(n->partials); cudaFree(n->grid_grad); cudaFree(n->loss_partials); cudaFree(n->grad_f32); cudaFree(n->loss_d);
    if (n->loss_pinned) cudaFreeHost(n->loss_pinned);
    cudaFree(n->loss4_d); n->loss4_d = nullptr;
    n->master = nullptr; n->params = nullptr; n->ema = nullptr; n->grads = nullptr; n->m1 = n->m2 = nullptr;
    n->steps = nullptr; n->partials = n->grid_grad = n->loss_partials = n->grad_f32 = n->loss_d = nullptr;
    n->loss_pinned = nullptr;
    for (int i = 0; i < Net::kHostChunks; ++i) {
        if (n->pin_in[i]) cudaFreeHost(n->pin_in[i]);
        if (n->pin_out[i]) cudaFreeHost(n->pin_out[i]);
        cudaFree(n->dev_in[i]); cudaFree(n->dev_out[i]);
        if (n->hstream[i]) cudaStreamDestroy(n->hstream[i]);
        if (n->hevent[i]) cudaEventDestroy(n->hevent[i]);
        n->pin_in[i] = n->pin_out[i] = n->dev_in[i] = n->dev_out[i] = nullptr;
        n->hstream[i] = nullptr; n->hevent[i] = nullptr;
    }
    n->host_chunk = 0;
    cudaFree(n->train_in_d); cudaFree(n->train_tgt_d);
    n->train_in_d = n->train_tgt_d = nullptr;
    if (n->err_host) cudaFreeHost(n->err_host);
    n->err_host = n->err_dev = nullptr;
}

// Uploads fp32 master weights, derives the fp16 copy (trainer.h:83-86), zeroes EMA (ema.h:93-94) and Adam state.
static int upload_master(Net* n, const float* host_master) {
    std::vector<__half> h(n->n_params);
    for (uint32_t i = 0; i < n->n_params; ++i) h[i] = __float2half_rn(host_master[i]);
    CU(cudaMemcpyAsync(n->master, host_master, (size_t)n->n_params * 4, cudaMemcpyHostToDevice, n->stream));
    CU(cudaMemcpyAsync(n->params, h.data(), (size_t)n->n_params * 2, cudaMemcpyHostToDevice, n->stream));
    CU(cudaMemsetAsync(n->ema, 0, (size_t)n->n_params * 2, n->stream));
    CU(cudaMemsetAsync(n->grads, 0, (size_t)n->n_params * 2, n->stream));
    CU(cudaMemsetAsync(n->m1, 0, (size_t)n->n_params * 4, n->stream));
    CU(cudaMemsetAsync(n->m2, 0, (size_t)n->n_params * 4, n->stream));
    CU(cudaMemsetAsync(n->steps, 0, (size_t)n->n_params * 4, n->stream));
    CU(cudaMemsetAsync(n->grad_f32, 0, (size_t)n->n_params * 4, n->stream));
    if (n->grid_grad) CU(cudaMemsetAsync(n->grid_grad, 0, (size_t)(n->n_params - n->n_mlp) * 4, n->stream));
    CU(cudaMemsetAsync(n->loss_d, 0, 4, n->stream));
    CU(launch_make_image(n->params, n->params_img, n->enc_width, n->n_mlp, n->stream));
    CU(cudaMemsetAsync(n->ema_img, 0, (size_t)n->n_mlp * 2, n->stream));
    CU(cudaStreamSynchronize(n->stream));  // host vectors go out of scope
    // The optimizer step counter restarts; the data-parallel exchange sequence (dp_seq: tags, flags, buffer parity) does
    // NOT -- stale tags of an earlier run can never match a later step (every rank must re-upload the same weights).
    n->step = 0;
    n->pending_tiles = 0;
    return NRC_B200_OK;
}

// Network::init_ (NRCNetwork.cu:106-112) -> create_from_config -> Trainer ctor -> initialize_params
// (trainer.h:68-95): xavier-uniform matrices (gpu_matrix.h:284-299, fully_fused_mlp.cu:866-891),
// grid U(-1e-4, 1e-4) (grid.h:1059-1062).  The random stream is splitmix64(seed), not tcnn's pcg32.
static int build_model(Net* n) {
    make_layout(n);
    n->lr = n->encoding == NRC_B200_FREQUENCY ? 1e-3f : 1e-2f;   // neural_radiance_caching.h:48-55
    n->eps = n->encoding == NRC_B200_FREQUENCY ? 1e-8f : 1e-15f; // NRCNetworkConfigs.h:96
    const size_t P = n->n_params;
    CU(cudaMalloc(&n->master, P * 4)); CU(cudaMalloc(&n->params, P * 2)); CU(cudaMalloc(&n->ema, P * 2));
    CU(cudaMalloc(&n->grads, P * 2)); CU(cudaMalloc(&n->m1, P * 4)); CU(cudaMalloc(&n->m2, P * 4));
    CU(cudaMalloc(&n->steps, P * 4)); CU(cudaMalloc(&n->grad_f32, P * 4));
    CU(cudaMalloc(&n->params_img, (size_t)n->n_mlp * 2));
    CU(cudaMalloc(&n->ema_img, (size_t)n->n_mlp * 2));
    CU(cudaMalloc(&n->barrier_counter, kBarrierWords * 4));
    CU(cudaMemset(n->barrier_counter, 0, kBarrierWords * 4));
    n->barrier_total = 0;
    CU(cudaMalloc(&n->partials, (size_t)kMaxTrainTiles * n->n_mlp * 4));
    CU(cudaMalloc(&n->loss_partials, kMaxTrainTiles * 4));
    CU(cudaMalloc(&n->loss_d, 4));
    if (P > n->n_mlp) CU(cudaMalloc(&n->grid_grad, (P - n->n_mlp) * 4));
    CU(cudaMallocHost(&n->loss_pinned, 16));
    CU(cudaHostAlloc(&n->err_host, 4, cudaHostAllocMapped));
    *n->err_host = 0;
    CU(cudaHostGetDevicePointer(&n->err_dev, n->err_host, 0));
    // cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute: set it for the device of THIS handle
    CU(configure_infer_kernels());
    CU(configure_train_kernels());
    CU(cudaMalloc(&n->loss4_d, 16));
    CU(cudaMemset(n->loss4_d, 0, 16));
    std::vector<float> m(P);
    uint64_t rng = n->seed;
    const int E = (int)n->enc_width;
    const int rows[6] = {64, 64, 64, 64, 64, 16}, cols[6] = {E, 64, 64, 64, 64, 64};
    size_t o = 0;
    for (int l = 0; l < 6; ++l) {
        const float scale = sqrtf(6.0f / (float)(rows[l] + cols[l]));
        for (int i = 0; i < rows[l] * cols[l]; ++i) m[o++] = u01(rng) * 2 * scale - scale;
    }
    for (; o < P; ++o) m[o] = u01(rng) * 2e-4f - 1e-4f;
    return upload_master(n, m.data());
}

// ---------------------------------------------------------------- lifecycle
int nrc_b200_create_ex(int encoding, int log2_hashmap_size, uint64_t seed, void* cu_stream, nrc_b200_handle* out) {
    if (!out) return fail(NRC_B200_ERR_INVALID, "nrc_b200_create: out is NULL");
    *out = nullptr;
    if (encoding != NRC_B200_FREQUENCY && encoding != NRC_B200_HASH)
        return fail(NRC_B200_ERR_INVALID, "Unsupported input encoding");  // NRCNetworkConfigs.h:129-131
    if (log2_hashmap_size <= 0) log2_hashmap_size = 15;
    if (log2_hashmap_size < 12 || log2_hashmap_size > 24)
        return fail(NRC_B200_ERR_INVALID, "log2_hashmap_size must be in [12, 24]");
    Net* n = new Net();
    n->encoding = encoding; n->log2_hashmap = log2_hashmap_size; n->seed = seed; n->stream = (cudaStream_t)cu_stream;
    if (const char* e = getenv("NRC_B200_DP_TIMEOUT_MS")) { const long v = atol(e); if (v > 0) n->dp_timeout_ms = (uint32_t)v; }
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDevice(&n->device);      // the device current at create owns the handle (Device.cpp:2296)
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, n->device);
    if (e != cudaSuccess) { delete n; return fail(NRC_B200_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e)); }
    if (prop.major != 10) {
        delete n;
        return fail(NRC_B200_ERR_CUDA, "nrc_b200 requires an sm_100a (B200) device; found sm_" +
                                           std::to_string(prop.major * 10 + prop.minor) + ". There is no fallback path.");
    }
    n->num_sms = prop.multiProcessorCount;
    int rc = build_model(n);
    if (rc != NRC_B200_OK) { free_all(n); delete n; return rc; }
    *out = n;
    return NRC_B200_OK;
}
int nrc_b200_create(int encoding, void* cu_stream, nrc_b200_handle* out) {
    return nrc_b200_create_ex(encoding, 15, 1337, cu_stream, out);
}
// Network::destroy (NRCNetwork.cu:35-39): frees every device / pinned allocation.  The (small) host struct stays behind as a
// tombstone so that later calls on the handle are the reference's silent no-ops (NRC_B200_ERR_DESTROYED) instead of a
// use-after-free; nrc_b200_reset revives it, nrc_b200_release frees it.
int nrc_b200_destroy(nrc_b200_handle h) {
    if (!h) return fail(NRC_B200_ERR_INVALID, "null handle");
    if (h->destroyed) return NRC_B200_ERR_DESTROYED;
    DeviceGuard guard_(h->device);
    cudaStreamSynchronize(h->stream);
    free_all(h);
    h->destroyed = true;
    return NRC_B200_OK;
}
int nrc_b200_release(nrc_b200_handle h) {
    if (!h) return NRC_B200_OK;
    if (!h->destroyed) nrc_b200_destroy(h);
    delete h;
    return NRC_B200_OK;
}
int nrc_b200_reset(nrc_b200_handle h, int encoding) {
    if (!h) return fail(NRC_B200_ERR_INVALID, "null handle");
    if (encoding != NRC_B200_FREQUENCY && encoding != NRC_B200_HASH)
        return fail(NRC_B200_ERR_INVALID, "Unsupported input encoding");
    DeviceGuard guard_(h->device);
    if (!h->destroyed) {
        CU(cudaStreamSynchronize(h->stream));
        free_all(h);
    }
    h->encoding = encoding;
    h->step = 0; h->dp_seq = 0; h->pending_tiles = 0;
    int rc = build_model(h);
    h->destroyed = rc != NRC_B200_OK;
    if (rc != NRC_B200_OK) free_all(h);
    return rc;
}
int nrc_b200_set_stream(nrc_b200_handle h, void* s) {
    if (!h) return fail(NRC_B200_ERR_INVALID, "null handle");
    if (h->destroyed) return NRC_B200_ERR_DESTROYED;
    h->stream = (cudaStream_t)s;
    return NRC_B200_OK;
}
int nrc_b200_set_learning_rate(nrc_b200_handle h, float lr) {
    if (!h) return fail(NRC_B200_ERR_INVALID, "null handle");
    if (h->destroyed) return NRC_B200_ERR_DESTROYED;
    h->lr = lr;
    return NRC_B200_OK;
}
int nrc_b200_get_learning_rate(nrc_b200_handle h, float* lr) {
    if (!h || !lr) return fail(NRC_B200_ERR_INVALID, "null argument");
    *lr = h->lr;                     // the hyper-parameter outlives destroy(), like Network::getLearningRate (NRCNetwork.cu:101-104)
    return NRC_B200_OK;
}
int nrc_b200_set_accumulator(nrc_b200_handle h, int mode) {
    if (!h) return fail(NRC_B200_ERR_INVALID, "null handle");
    if (h->destroyed) return NRC_B200_ERR_DESTROYED;
    if (mode != NRC_B200_ACC_FP32 && mode != NRC_B200_ACC_FP16) return fail(NRC_B200_ERR_INVALID, "unknown accumulator mode");
    h->acc16 = mode == NRC_B200_ACC_FP16;
    return NRC_B200_OK;
}
int nrc_b200_launch_count(nrc_b200_handle h, uint64_t* c) {
    if (!h || !c) return fail(NRC_B200_ERR_INVALID, "null argument");
    *c = h->launches;
    return NRC_B200_OK;
}

// ---------------------------------------------------------------- parameters
int nrc_b200_n_params(nrc_b200_handle h, uint32_t* n_params, uint32_t* n_mlp) {
    if (!h) return fail(NRC_B200_ERR_INVALID, "null handle");
    if (n_params) *n_params = h->n_params;
    if (n_mlp) *n_mlp = h->n_mlp;
    return NRC_B200_OK;
}
static void* param_ptr(Net* n, int kind, size_t* elem) {
    switch (kind) {
        case NRC_B200_PARAMS_MASTER_F32: *elem = 4; return n->master;
        case NRC_B200_PARAMS_TRAIN_F16: *elem = 2; return n->params;
        case NRC_B200_PARAMS_EMA_F16: *elem = 2; return n->ema;
        case NRC_B200_PARAMS_GRAD_F16: *elem = 2; return n->grads;
        case NRC_B200_ADAM_M1_F32: *elem = 4; return n->m1;
        case NRC_B200_ADAM_M2_F32: *elem = 4; return n->m2;
        case NRC_B200_GRAD_F32: *elem = 4; return n->grad_f32;
    }
    return nullptr;
}
int nrc_b200_get_params(nrc_b200_handle h, int kind, void* dst) {
    if (!h || !dst || h->destroyed) return fail(NRC_B200_ERR_INVALID, "bad handle/argument");
    DeviceGuard guard_(h->device);
    size_t es; void* p = param_ptr(h, kind, &es);
    if (!p) return fail(NRC_B200_ERR_INVALID, "unknown parameter kind");
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaMemcpy(dst, p, es * h->n_params, cudaMemcpyDeviceToHost));
    return NRC_B200_OK;
}
int nrc_b200_set_params(nrc_b200_handle h, int kind, const void* src) {
    if (!h || !src || h->destroyed) return fail(NRC_B200_ERR_INVALID, "bad handle/argument");
    DeviceGuard guard_(h->device);
    size_t es; void* p = param_ptr(h, kind, &es);
    if (!p) return fail(NRC_B200_ERR_INVALID, "unknown parameter kind");
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaMemcpy(p, src, es * h->n_params, cudaMemcpyHostToDevice));
    if (kind == NRC_B200_PARAMS_TRAIN_F16 || kind == NRC_B200_PARAMS_EMA_F16) {   // keep the UMMA images current
        CU(launch_make_image(kind == NRC_B200_PARAMS_TRAIN_F16 ? h->params : h->ema,
                             kind == NRC_B200_PARAMS_TRAIN_F16 ? h->params_img : h->ema_img, h->enc_width, h->n_mlp, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    return NRC_B200_OK;
}
int nrc_b200_load_master(nrc_b200_handle h, const float* m) {
    if (!h || !m || h->destroyed) return fail(NRC_B200_ERR_INVALID, "bad handle/argument");
    DeviceGuard guard_(h->device);
    return upload_master(h, m);
}

// ---------------------------------------------------------------- inference
static cudaError_t run_infer(Net* h, const float* q, float* r, uint32_t n, cudaStream_t stream, const AccumArgs* acc = nullptr) {
    return launch_infer(h->encoding, h->acc16, q, r, n, h->ema, h->ema_img, h->gm, h->num_sms, stream, acc);
}
int nrc_b200_infer(nrc_b200_handle h, const float* q, float* r, uint32_t n, void* cu_stream) {
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (n == 0) return NRC_B200_OK;
    if (!q || !r) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    CU(run_infer(h, q, r, n, h->stream));
    h->launches += 1;
    return NRC_B200_OK;
}

static bool is_pinned_host(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// Host buffers: 3-deep pipeline of (pinned copy-in -> H2D -> kernel -> D2H -> copy-out) over chunks.
int nrc_b200_infer_host(nrc_b200_handle h, const float* q, float* r, uint32_t n) {
    ENTER(h);
    if (n == 0) return NRC_B200_OK;
    if (!q || !r) return fail(NRC_B200_ERR_INVALID, "null host buffer");
    const uint32_t want = 1u << 18;  // queries per chunk: 15.7 MB in, 3.1 MB out
    if (h->host_chunk < want) {
        for (int i = 0; i < Net::kHostChunks; ++i) {
            CU(cudaMallocHost(&h->pin_in[i], (size_t)want * 60)); CU(cudaMallocHost(&h->pin_out[i], (size_t)want * 12));
            CU(cudaMalloc(&h->dev_in[i], (size_t)want * 60)); CU(cudaMalloc(&h->dev_out[i], (size_t)want * 12));
            CU(cudaStreamCreateWithFlags(&h->hstream[i], cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&h->hevent[i], cudaEventDisableTiming));
        }
        h->host_chunk = want;
    }
    CU(cudaStreamSynchronize(h->stream));  // weights written by earlier training on h->stream
    // Page-locked caller buffers (cudaHostAlloc / cudaHostRegister) are DMA'd directly; pageable ones go through
    // the pinned staging ring.
    const bool q_pinned = is_pinned_host(q), r_pinned = is_pinned_host(r);
    const uint32_t chunk = h->host_chunk;
    const uint32_t n_chunks = (n + chunk - 1) / chunk;
    for (uint32_t c = 0; c < n_chunks + Net::kHostChunks; ++c) {
        if (c >= Net::kHostChunks) {  // retire chunk c - kHostChunks
            const uint32_t d = c - Net::kHostChunks, s = d % Net::kHostChunks;
            if (d < n_chunks) {
                CU(cudaEventSynchronize(h->hevent[s]));
                const uint32_t cnt = std::min(chunk, n - d * chunk);
                if (!r_pinned) memcpy(r + (size_t)d * chunk * 3, h->pin_out[s], (size_t)cnt * 12);
            }
        }
        if (c < n_chunks) {
            const uint32_t s = c % Net::kHostChunks, cnt = std::min(chunk, n - c * chunk);
            const float* src = q + (size_t)c * chunk * 15;
            if (!q_pinned) { memcpy(h->pin_in[s], src, (size_t)cnt * 60); src = h->pin_in[s]; }
            CU(cudaMemcpyAsync(h->dev_in[s], src, (size_t)cnt * 60, cudaMemcpyHostToDevice, h->hstream[s]));
            CU(run_infer(h, h->dev_in[s], h->dev_out[s], cnt, h->hstream[s]));
            h->launches += 1;
            float* dst = r_pinned ? r + (size_t)c * chunk * 3 : h->pin_out[s];
            CU(cudaMemcpyAsync(dst, h->dev_out[s], (size_t)cnt * 12, cudaMemcpyDeviceToHost, h->hstream[s]));
            CU(cudaEventRecord(h->hevent[s], h->hstream[s]));
        }
    }
    return NRC_B200_OK;
}

// ---------------------------------------------------------------- training
// `first_step` is the (first) optimizer step this launch applies; n_steps > 1: steps first_step .. first_step + n_steps - 1.
// Nothing in the handle is advanced here: the callers commit step / barrier / exchange counters only AFTER the launch has
// been accepted (a failed launch must not leave the host's barrier target ahead of the device counter).
static TrainArgs train_args(Net* h, const float* q, const float* t, uint32_t n_local, uint32_t n_global, int apply,
                            uint32_t n_steps = 1) {
    TrainArgs a{};
    const uint32_t first_step = h->step + 1;
    a.in = q; a.targets = t; a.n_local = n_local; a.n_global = n_global;
    a.master = h->master; a.params = h->params; a.params_img = h->params_img; a.ema = h->ema; a.ema_img = h->ema_img; a.grads = h->grads;
    a.m1 = h->m1; a.m2 = h->m2; a.steps = h->steps;
    a.partials = h->partials; a.grid_grad = h->grid_grad; a.loss_partials = h->loss_partials;
    a.grad_f32 = h->grad_f32; a.loss_out = h->loss_d;
    a.barrier_counter = h->barrier_counter;
    a.wide_grid = (h->encoding == NRC_B200_HASH && h->num_sms >= (int)kWideGrid) ? 1 : 0;
    a.barrier_target = h->barrier_total + train_barrier_arrivals(n_local, a.wide_grid);   // top counter once every CTA reached the FIRST barrier
    a.n_params = h->n_params; a.n_mlp = h->n_mlp;
    a.lr = h->lr; a.beta1 = h->beta1; a.beta2 = h->beta2; a.eps = h->eps; a.l2_reg = h->l2_reg;
    a.ema_decay = h->ema_decay; a.loss_scale = h->loss_scale;
    {   // x / 2^k == x * 2^-k exactly (no denormals in reach: fp16 gradients, scale 128): the optimizer multiplies then
        int ex = 0;
        a.loss_scale_pow2 = (h->loss_scale > 0.0f && frexpf(h->loss_scale, &ex) == 0.5f) ? 1 : 0;
        a.inv_loss_scale = 1.0f / h->loss_scale;
    }
    // EmaOptimizer::step (ema.h:107-108): std::pow(float, unsigned) promotes to double
    a.n_steps = n_steps;
    for (uint32_t s = 0; s < n_steps && s < 4; ++s) {
        a.ema_debias_old[s] = 1.0f - (float)pow((double)h->ema_decay, (double)(first_step + s - 1));
        a.ema_debias_new[s] = 1.0f / (1.0f - (float)pow((double)h->ema_decay, (double)(first_step + s)));
    }
    a.apply = apply;
    // "reference arithmetic" option: fp16 accumulators in the query kernel AND in the forward / backward chain of the training
    // step (fully_fused_mlp.cu:68,150-259); the Hash grid gradients are scattered as half products with half2 atomics
    // (grid.h:252-255,301-319).  Weight gradients stay fp32 sums (the reference: fp16 split-K); the data-parallel exchange
    // always carries fp32 sums.
    a.acc16 = h->acc16;
    a.grid_grad_fp16 = (h->acc16 && h->encoding == NRC_B200_HASH && h->dp_world <= 1) ? 1 : 0;
    a.timeline = h->timeline;
    a.perm_n = 0; a.perm_seed = 0; a.perm_base = 0;
    a.opt_step = first_step;
    a.dp_world = 1; a.dp_rank = 0; a.dp_seq = h->dp_seq + 1; a.x_stride = h->x_stride;
    a.error_flag = h->err_dev; a.dp_timeout_ms = h->dp_timeout_ms;
    if (apply && h->dp_world > 1) {
        a.dp_world = (uint32_t)h->dp_world; a.dp_rank = (uint32_t)h->dp_rank;
        for (int r = 0; r < h->dp_world; ++r) {
            a.xbuf[r] = (float*)h->dp_peer[r];
            a.xflag[r] = (uint32_t*)((char*)h->dp_peer[r] + h->xbuf_bytes);
        }
    }
    return a;
}
// Launches a step kernel and, on success, commits what it consumed: optimizer steps, grid-barrier arrivals, exchange steps.
static int launch_and_commit(Net* h, const TrainArgs& a, uint32_t n_barriers) {
    CU(launch_train_step(h->encoding, a, h->gm, h->stream));
    h->launches += 1;
    h->barrier_total += n_barriers * train_barrier_arrivals(a.n_local, a.wide_grid);
    if (a.apply) {
        h->step += a.n_steps;
        if (a.dp_world > 1) h->dp_seq += a.n_steps;
        h->pending_tiles = 0;
    }
    return NRC_B200_OK;
}
// Samples the loss of one step is normalised by.  A handle connected to data-parallel peers trains on the GLOBAL batch of
// dp_world local batches whichever entry point is used (every rank must make the same calls).
static uint32_t global_batch(const Net* h, uint32_t n_local) { return h->dp_world > 1 ? n_local * (uint32_t)h->dp_world : n_local; }
static int read_loss(Net* h, const float* src_d, int n, float* loss_h) {   // Trainer::loss: D2H + stream sync (reduce_sum.h:146-149)
    CU(cudaMemcpyAsync(h->loss_pinned, src_d, 4 * n, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (*(volatile uint32_t*)h->err_host) return device_error(h);
    float l = 0.0f;
    for (int b = 0; b < n; ++b) l += h->loss_pinned[b];
    *loss_h = l / (float)n;
    return NRC_B200_OK;
}

int nrc_b200_train_grad(nrc_b200_handle h, const float* q, const float* t, uint32_t n_local, uint32_t n_global,
                        float** grad_d, void* cu_stream) {
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (!q || !t) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    if (n_local == 0 || n_local > NRC_B200_BATCH_SIZE || n_local % 128 != 0 || n_global < n_local)
        return fail(NRC_B200_ERR_INVALID, "n_local must be a positive multiple of 128, <= 16384 and <= n_global");
    int rc = launch_and_commit(h, train_args(h, q, t, n_local, n_global, /*apply=*/0), 1);
    if (rc != NRC_B200_OK) return rc;
    h->pending_tiles = n_local / 128;
    if (grad_d) *grad_d = h->grad_f32;
    return NRC_B200_OK;
}

int nrc_b200_train_apply(nrc_b200_handle h, void* cu_stream, float* loss_h) {
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (h->pending_tiles == 0) return fail(NRC_B200_ERR_INVALID, "train_apply without a pending gradient");
    TrainArgs a = train_args(h, nullptr, nullptr, 0, 0, 1);
    a.dp_world = 1;                  // the caller has all-reduced grad_f32 itself
    CU(launch_adam_ema(a, h->stream));
    h->launches += 1;
    h->step += 1;
    h->pending_tiles = 0;
    if (loss_h) return read_loss(h, h->loss_d, 1, loss_h);
    return NRC_B200_OK;
}

int nrc_b200_train(nrc_b200_handle h, const float* q, const float* t, void* cu_stream, float* loss_h) {
    ENTER(h);                        // NRCNetwork.cu:43
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (!q || !t) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    // one cooperative launch: forward + loss + backward + gradient reduction (+ all-reduce over the peers) + Adam + EMA
    int rc = launch_and_commit(h, train_args(h, q, t, NRC_B200_BATCH_SIZE, global_batch(h, NRC_B200_BATCH_SIZE), /*apply=*/1), 1);
    if (rc != NRC_B200_OK) return rc;
    if (loss_h) return read_loss(h, h->loss_d, 1, loss_h);
    return NRC_B200_OK;
}

// Host-buffer training step: H2D of one batch (983 KB + 197 KB) on the training stream, then the same kernels.
int nrc_b200_train_host(nrc_b200_handle h, const float* q_h, const float* t_h, float* loss_h) {
    ENTER(h);
    if (!q_h || !t_h) return fail(NRC_B200_ERR_INVALID, "null host buffer");
    if (!h->train_in_d) {
        CU(cudaMalloc(&h->train_in_d, (size_t)NRC_B200_BATCH_SIZE * 60));
        CU(cudaMalloc(&h->train_tgt_d, (size_t)NRC_B200_BATCH_SIZE * 12));
    }
    CU(cudaMemcpyAsync(h->train_in_d, q_h, (size_t)NRC_B200_BATCH_SIZE * 60, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(h->train_tgt_d, t_h, (size_t)NRC_B200_BATCH_SIZE * 12, cudaMemcpyHostToDevice, h->stream));
    return nrc_b200_train(h, h->train_in_d, h->train_tgt_d, nullptr, loss_h);
}

// ---------------------------------------------------------------- data-parallel training over NVLink peer memory
// Step 1 (every rank): allocate the exchange buffer and export its CUDA IPC handle (64 bytes).
static int dp_alloc(Net* h) {
    CU(cudaStreamSynchronize(h->stream));
    dp_teardown(h);
    // per (parity, source rank) row: [(n_mlp + 1 -> 32) tagged float2 slots | n_grid floats]  (nrc_train.cu phase 3)
    h->x_stride = 2 * ((h->n_mlp + 1 + 31) / 32 * 32) + ((h->n_params - h->n_mlp) + 31) / 32 * 32;
    h->xbuf_bytes = (size_t)2 * kMaxDpRanks * h->x_stride * sizeof(float);
    const size_t flag_bytes = (size_t)2 * kMaxDpRanks * kDpFlagSlots * sizeof(uint32_t);
    CU(cudaMalloc(&h->dp_local, h->xbuf_bytes + flag_bytes));
    CU(cudaMemset(h->dp_local, 0, h->xbuf_bytes + flag_bytes));
    CU(cudaDeviceSynchronize());
    h->dp_seq = 0;                   // tags / flags start at 1 on a zeroed buffer; never reset while connected (weights may be re-uploaded)
    return NRC_B200_OK;
}
int nrc_b200_dp_export(nrc_b200_handle h, void* ipc_handle_out) {
    if (!h || h->destroyed || !ipc_handle_out) return fail(NRC_B200_ERR_INVALID, "bad handle/argument");
    DeviceGuard guard_(h->device);
    int rc = dp_alloc(h);
    if (rc != NRC_B200_OK) return rc;
    cudaIpcMemHandle_t mh;
    CU(cudaIpcGetMemHandle(&mh, h->dp_local));
    static_assert(sizeof(mh) == NRC_B200_IPC_HANDLE_BYTES, "IPC handle size");
    memcpy(ipc_handle_out, &mh, sizeof(mh));
    return NRC_B200_OK;
}
// Step 2 (every rank, after an all-gather of the handles): map the peers' buffers.  The optimizer state must be
// identical on all ranks (same seed / same nrc_b200_load_master) and the step counters equal.
int nrc_b200_dp_connect(nrc_b200_handle h, int rank, int world, const void* ipc_handles) {
    if (!h || h->destroyed || !ipc_handles) return fail(NRC_B200_ERR_INVALID, "bad handle/argument");
    if (world < 1 || world > kMaxDpRanks || rank < 0 || rank >= world) return fail(NRC_B200_ERR_INVALID, "bad rank/world");
    if (!h->dp_local) return fail(NRC_B200_ERR_INVALID, "nrc_b200_dp_export must be called first");
    DeviceGuard guard_(h->device);
    for (int r = 0; r < world; ++r) {
        if (r == rank) { h->dp_peer[r] = h->dp_local; continue; }
        cudaIpcMemHandle_t mh;
        memcpy(&mh, (const char*)ipc_handles + (size_t)r * sizeof(mh), sizeof(mh));
        CU(cudaIpcOpenMemHandle(&h->dp_peer[r], mh, cudaIpcMemLazyEnablePeerAccess));
    }
    h->dp_rank = rank; h->dp_world = world;
    return NRC_B200_OK;
}
// Same exchange between handles of ONE process (one handle per GPU, driven from one or several host threads): peer
// access instead of CUDA IPC.  handles[r] is rank r's handle; allocates every exchange buffer and connects all of them.
int nrc_b200_dp_connect_local(nrc_b200_handle* handles, int world) {
    if (!handles || world < 1 || world > kMaxDpRanks) return fail(NRC_B200_ERR_INVALID, "bad handles/world");
    for (int r = 0; r < world; ++r) {
        Net* h = handles[r];
        if (!h || h->destroyed) return fail(NRC_B200_ERR_INVALID, "bad handle");
        if (h->n_params != handles[0]->n_params) return fail(NRC_B200_ERR_INVALID, "handles must share one configuration");
        DeviceGuard guard_(h->device);
        int rc = dp_alloc(h);
        if (rc != NRC_B200_OK) return rc;
        for (int p = 0; p < world; ++p) {
            if (handles[p]->device == h->device) continue;
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, h->device, handles[p]->device));
            if (!can) return fail(NRC_B200_ERR_CUDA, "no peer access between the devices of the handles");
            cudaError_t e = cudaDeviceEnablePeerAccess(handles[p]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU(e);
            cudaGetLastError();
        }
    }
    for (int r = 0; r < world; ++r) {
        Net* h = handles[r];
        for (int p = 0; p < world; ++p) h->dp_peer[p] = handles[p]->dp_local;
        h->dp_rank = r; h->dp_world = world; h->dp_shared_process = true;
    }
    return NRC_B200_OK;
}
// One optimizer step on a global batch of n_global samples of which this rank holds n_local: ONE cooperative kernel
// per rank does forward/backward, the gradient all-reduce (peer stores + flags over NVLink) and Adam + EMA.
// Every rank must call it for every step; loss_h (optional, blocks) receives the global loss.
int nrc_b200_train_dp(nrc_b200_handle h, const float* q, const float* t, uint32_t n_local, uint32_t n_global,
                      void* cu_stream, float* loss_h) {
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (!q || !t) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    if (n_local == 0 || n_local > NRC_B200_BATCH_SIZE || n_local % 128 != 0 || n_global < n_local)
        return fail(NRC_B200_ERR_INVALID, "n_local must be a positive multiple of 128, <= 16384 and <= n_global");
    int rc = launch_and_commit(h, train_args(h, q, t, n_local, n_global, /*apply=*/1), 1);
    if (rc != NRC_B200_OK) return rc;
    if (loss_h) return read_loss(h, h->loss_d, 1, loss_h);
    return NRC_B200_OK;
}

// ---------------------------------------------------------------- NEXT rows: the per-frame sequence around the network
int nrc_b200_accumulate_render_radiance(const float* radiance_d, const float* throughput_d, float* output_rgba_d,
                                        uint32_t n_pixels, int render_mode, uint32_t iteration_index, void* cu_stream) {
    if (n_pixels && (!radiance_d || !throughput_d || !output_rgba_d)) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    CU(launch_accumulate(radiance_d, throughput_d, output_rgba_d, n_pixels, render_mode, iteration_index, (cudaStream_t)cu_stream));
    return NRC_B200_OK;
}
int nrc_b200_propagate_train_radiance(const nrc_b200_suffix_end_vertex* end_vertices_d, const float* end_radiance_d,
                                      const nrc_b200_training_record* records_d, float* targets_d, uint32_t n_tiles,
                                      void* cu_stream) {
    if (n_tiles && (!end_vertices_d || !end_radiance_d || !records_d || !targets_d)) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    static_assert(sizeof(nrc_b200_training_record) == 28 && sizeof(nrc_b200_suffix_end_vertex) == 16, "record layouts");
    CU(launch_propagate(end_vertices_d, end_radiance_d, records_d, targets_d, n_tiles, (cudaStream_t)cu_stream));
    return NRC_B200_OK;
}
uint32_t nrc_b200_permutation(uint32_t i, uint32_t seed) { return permute16(i & 0xFFFFu, seed); }

// n_batches consecutive optimizer steps on q + b*16384*15 / t + b*16384*3 (or, perm_n > 0, on the fused shuffle of the
// frame's records), losses into loss4_d[b].  Frequency: ONE launch (MULTI instantiation, single GPU and data parallel);
// Hash: one launch per batch (the grid scatter makes a step 4x longer and the in-kernel barriers cost more than the
// launch gaps: 90 vs 76 us per step measured).
static int train_steps(Net* h, const float* q, const float* t, int n_batches, uint32_t perm_n, uint32_t perm_seed) {
    const uint32_t n_global = global_batch(h, NRC_B200_BATCH_SIZE);
    // Hash: NRC_B200_HASH_MULTI=0 falls back to one launch per batch (the round-1 form, kept for A/B measurements)
    static const bool hash_multi = [] { const char* e = getenv("NRC_B200_HASH_MULTI"); return !(e && e[0] == '0'); }();
    if (h->encoding == NRC_B200_HASH && !hash_multi) {
        for (int b = 0; b < n_batches; ++b) {
            TrainArgs a = perm_n ? train_args(h, q, t, NRC_B200_BATCH_SIZE, n_global, 1)
                                 : train_args(h, q + (size_t)b * NRC_B200_BATCH_SIZE * 15, t + (size_t)b * NRC_B200_BATCH_SIZE * 3,
                                              NRC_B200_BATCH_SIZE, n_global, 1);
            a.perm_n = perm_n; a.perm_seed = perm_seed; a.perm_base = (uint32_t)b * NRC_B200_BATCH_SIZE;
            a.loss_out = h->loss4_d + b;
            int rc = launch_and_commit(h, a, 1);
            if (rc != NRC_B200_OK) return rc;
        }
        return NRC_B200_OK;
    }
    TrainArgs a = train_args(h, q, t, NRC_B200_BATCH_SIZE, n_global, 1, (uint32_t)n_batches);
    a.perm_n = perm_n; a.perm_seed = perm_seed; a.perm_base = 0;
    a.loss_out = h->loss4_d;
    return launch_and_commit(h, a, 2 * (uint32_t)n_batches - 1);
}
// Device::nrcTrainRadiance (Device.cpp:1482-1513): NUM_BATCHES consecutive train() calls.
int nrc_b200_train_batches(nrc_b200_handle h, const float* q, const float* t, int n_batches, void* cu_stream, float* loss_h) {
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (!q || !t) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    if (n_batches < 1 || n_batches > NRC_B200_NUM_BATCHES) return fail(NRC_B200_ERR_INVALID, "n_batches must be in [1, 4]");
    int rc = train_steps(h, q, t, n_batches, 0, 0);
    if (rc != NRC_B200_OK) return rc;
    if (loss_h) return read_loss(h, h->loss4_d, n_batches, loss_h);   // mean of the per-batch losses, as Device::nrcTrainRadiance reports it
    return NRC_B200_OK;
}
int nrc_b200_permute_train_data(const float* q_src_d, const float* t_src_d, float* q_dst_d, float* t_dst_d,
                                int num_records, uint32_t seed, void* cu_stream) {
    if (!q_src_d || !t_src_d || !q_dst_d || !t_dst_d) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    const int n = std::min(num_records, NRC_B200_BATCH_SIZE * NRC_B200_NUM_BATCHES);
    if (n <= 0) return NRC_B200_OK;        // nrc_helpers.cu:238: nothing to do
    CU(launch_permute(q_src_d, t_src_d, q_dst_d, t_dst_d, (uint32_t)n, seed, (cudaStream_t)cu_stream));
    return NRC_B200_OK;
}
// permute_train_data with a caller-supplied permutation (the reference's cuRAND + CUB result, nrc_helpers.cu:226-249).
int nrc_b200_permute_train_data_with(const float* q_src_d, const float* t_src_d, float* q_dst_d, float* t_dst_d,
                                     const int* permutation_d, int num_records, void* cu_stream) {
    if (!q_src_d || !t_src_d || !q_dst_d || !t_dst_d || !permutation_d) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    const int n = std::min(num_records, NRC_B200_BATCH_SIZE * NRC_B200_NUM_BATCHES);
    if (n <= 0) return NRC_B200_OK;
    CU(launch_permute_with(q_src_d, t_src_d, q_dst_d, t_dst_d, permutation_d, (uint32_t)n, (cudaStream_t)cu_stream));
    return NRC_B200_OK;
}
void nrc_b200_adjust_tile_size(int tile_x, int tile_y, int num_train_records, int* new_x, int* new_y) {
    // Device::adjustTileSize, Device.cpp:818-828
    const float ratio = (float)(num_train_records * 1.25f) / (float)(NRC_B200_BATCH_SIZE * NRC_B200_NUM_BATCHES);
    const float r = sqrtf(ratio);
    if (new_x) *new_x = std::max((int)(tile_x * r + 0.5f), 2);
    if (new_y) *new_y = std::max((int)(tile_y * r + 0.5f), 2);
}

// permute_train_data + Network::train on batch `batch_index` in one launch: the training kernel's loader gathers
// src[permutation(d) % num_records] itself.
int nrc_b200_train_shuffled(nrc_b200_handle h, const float* q_src_d, const float* t_src_d, int num_records, uint32_t seed,
                            uint32_t batch_index, void* cu_stream, float* loss_h) {
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    if (!q_src_d || !t_src_d) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    if (batch_index >= NRC_B200_NUM_BATCHES) return fail(NRC_B200_ERR_INVALID, "batch_index must be < 4");
    const int n = std::min(num_records, NRC_B200_BATCH_SIZE * NRC_B200_NUM_BATCHES);
    if (n <= 0) return NRC_B200_OK;
    TrainArgs a = train_args(h, q_src_d, t_src_d, NRC_B200_BATCH_SIZE, global_batch(h, NRC_B200_BATCH_SIZE), /*apply=*/1);
    a.perm_n = (uint32_t)n; a.perm_seed = seed; a.perm_base = batch_index * NRC_B200_BATCH_SIZE;
    int rc = launch_and_commit(h, a, 1);
    if (rc != NRC_B200_OK) return rc;
    if (loss_h) return read_loss(h, h->loss_d, 1, loss_h);
    return NRC_B200_OK;
}

// The NRC part of Device::render (Device.cpp:2495-2515) in one call: infer -> accumulate -> propagate -> (fused shuffle) ->
// 4 x train.  No host synchronisation unless stat != NULL (then one read-back of the four losses at the end; the
// reference synchronises after every training step).
int nrc_b200_update(nrc_b200_handle h, const nrc_b200_frame* f, void* cu_stream, nrc_b200_training_stat* stat) {
    if (!f) return fail(NRC_B200_ERR_INVALID, "null argument");
    ENTER(h);
    if (cu_stream) h->stream = (cudaStream_t)cu_stream;
    const bool skip_render_cache = f->render_mode == 1 || f->render_mode == 3;   // NoCache, CacheFirstVertex
    const float* q = f->queries_d;
    float* r = f->results_d;
    uint32_t n = f->num_pixels + f->num_tiles;
    if (skip_render_cache) { q += (size_t)f->num_pixels * 15; r += (size_t)f->num_pixels * 3; n = f->num_tiles; }
    if (n && (!q || !r)) return fail(NRC_B200_ERR_INVALID, "null device buffer");
    // nrcInferRadiance + nrcAccumulateRadiance in ONE launch: the query kernel's epilogue accumulates the pixel part
    AccumArgs acc{nullptr, nullptr, 0, 0, 0.0f};
    if (!skip_render_cache && f->throughput_d && f->output_rgba_d)
        acc = AccumArgs{f->throughput_d, reinterpret_cast<float4*>(f->output_rgba_d), f->num_pixels, f->render_mode,
                        1.0f / (float)(f->iteration_index + 1)};
    CU(run_infer(h, q, r, n, h->stream, &acc));
    h->launches += 1;
    const int n_rec = std::min(f->num_train_records, NRC_B200_BATCH_SIZE * NRC_B200_NUM_BATCHES);
    if (stat) { stat->loss = 0.0f; stat->num_train_records = n_rec; }
    if (n_rec > 0) {
        if (!f->train_queries_d || !f->train_targets_d) return fail(NRC_B200_ERR_INVALID, "null training buffer");
        if (f->end_vertices_d && f->records_d && f->num_tiles) {                   // nrcPropagateRadiance
            CU(launch_propagate(f->end_vertices_d, f->results_d + (size_t)f->num_pixels * 3, f->records_d, f->train_targets_d,
                                f->num_tiles, h->stream));
            h->launches += 1;
        }
        // nrcShuffleTrainingData + nrcTrainRadiance: the four steps with the shuffle fused into their loaders
        int rc = train_steps(h, f->train_queries_d, f->train_targets_d, NRC_B200_NUM_BATCHES, (uint32_t)n_rec, f->shuffle_seed);
        if (rc != NRC_B200_OK) return rc;
        if (stat) return read_loss(h, h->loss4_d, NRC_B200_NUM_BATCHES, &stat->loss);
    }
    return NRC_B200_OK;
}

// Debug: per-tile clock64 stamps of the last training step [128 tiles][8 slots]:
// 0 start, 1 encoded, 2 forward done, 3 backward + weight gradients done, 4 partials stored, 5 past the grid barrier, 6 end.
int nrc_b200_debug_train_timeline(nrc_b200_handle h, int enable, long long* dst_host) {
    if (!h || h->destroyed) return fail(NRC_B200_ERR_INVALID, "bad handle");
    DeviceGuard guard_(h->device);
    if (enable && !h->timeline) {
        CU(cudaMalloc(&h->timeline, 148 * 8 * sizeof(long long)));
        CU(cudaMemset(h->timeline, 0, 148 * 8 * sizeof(long long)));
    }
    if (dst_host && h->timeline) {
        CU(cudaStreamSynchronize(h->stream));
        CU(cudaMemcpy(dst_host, h->timeline, 128 * 8 * sizeof(long long), cudaMemcpyDeviceToHost));
    }
    if (!enable && h->timeline) { CU(cudaStreamSynchronize(h->stream)); cudaFree(h->timeline); h->timeline = nullptr; }
    return NRC_B200_OK;
}

// Measurement aid for bench.py: read bandwidth (GB/s) the SMs get from an L2-resident buffer of `bytes` on the current device.
int nrc_b200_debug_l2_read_bandwidth(uint64_t bytes, int reps, float* gbs) {
    if (!gbs || bytes < (1u << 20) || reps < 1) return fail(NRC_B200_ERR_INVALID, "bad argument");
    CU(l2_read_probe((size_t)bytes & ~(size_t)15, reps, gbs));
    return NRC_B200_OK;
}

int nrc_b200_last_loss_device(nrc_b200_handle h, const float** loss_d) {
    if (!h || !loss_d || h->destroyed) return fail(NRC_B200_ERR_INVALID, "bad handle/argument");
    *loss_d = h->loss_d;
    return NRC_B200_OK;
}

#include <mutex>
#include "nrc_noalloc.inc"   // nrc_b200_train_grad_noalloc: static device scratch, no allocation
