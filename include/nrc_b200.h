/* nrc_b200.h -- C ABI of the B200-native neural radiance cache.
 *
 * Drop-in boundary for the reference's cache interface `class nrc::Network`
 * (/root/reference/nrc/inc/NRCNetwork.h:20-75, implemented in nrc/src/NRCNetwork.cu:22-128).
 * Plain pointers and sizes only; every function returns 0 on success or a negative nrc_b200_status and
 * leaves a message for nrc_b200_last_error().  The C++ shim include/NRCNetwork_b200.h re-creates
 * `nrc::Network` on top of these calls with the reference's exact signatures and exception behaviour.
 *
 * Data contract (unchanged from the reference, nrc/src/NRCNetwork.cu:41-83, nrc/src/Device.cpp:198,1272-1288):
 *   queries   : device pointer, `n` records of 15 float32 (nrc::RadianceQuery, 60 B, sample-major)
 *   radiance  : device pointer, `n` records of 3 float32
 *   train     : exactly NRC_B200_BATCH_SIZE (16384) queries + targets per call
 * Unlike the reference, rows in [n, round_up(n,256)) of the caller's buffers are never touched.
 */
#ifndef NRC_B200_H
#define NRC_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define NRC_B200_INPUT_DIMS 15       /* nrc::NN_INPUT_DIMS,  neural_radiance_caching.h:35-43 */
#define NRC_B200_OUTPUT_DIMS 3       /* nrc::NN_OUTPUT_DIMS, neural_radiance_caching.h:44    */
#define NRC_B200_BATCH_SIZE 16384    /* nrc::BATCH_SIZE,     neural_radiance_caching.h:31    */
#define NRC_B200_NUM_BATCHES 4       /* nrc::NUM_BATCHES,    neural_radiance_caching.h:29    */

typedef enum nrc_b200_encoding {     /* nrc::InputEncoding, neural_radiance_caching.h:24-27 */
    NRC_B200_FREQUENCY = 0,
    NRC_B200_HASH = 1
} nrc_b200_encoding;

typedef enum nrc_b200_status {
    NRC_B200_OK = 0,
    NRC_B200_ERR_INVALID = -1,   /* bad argument (reference: std::invalid_argument, NRCNetworkConfigs.h:129-131) */
    NRC_B200_ERR_CUDA = -2,      /* CUDA runtime failure (reference: std::runtime_error from CUDA_CHECK_THROW) */
    NRC_B200_ERR_DESTROYED = -3  /* informational: call on a destroyed (not yet released) handle is a no-op (NRCNetwork.cu:43,66) */
} nrc_b200_status;

typedef struct nrc_b200_network* nrc_b200_handle;

/* Network::init<>(stream, encoding)  NRCNetwork.h:26-31, NRCNetwork.cu:106-112.
 * Creates model + optimizer state on the current CUDA device; `cu_stream` is a CUstream / cudaStream_t.
 * `log2_hashmap_size` <= 0 selects the reference's 15 (NRCNetworkConfigs.h:101); `seed` seeds the
 * xavier-uniform initialisation (the reference uses 1337, trainer.h:50-57). */
int nrc_b200_create(int encoding, void* cu_stream, nrc_b200_handle* out);
int nrc_b200_create_ex(int encoding, int log2_hashmap_size, uint64_t seed, void* cu_stream, nrc_b200_handle* out);
/* Network::destroy()  NRCNetwork.cu:35-39.  Frees every device / pinned allocation.  The handle itself stays valid as a
 * tombstone: later calls on it return NRC_B200_ERR_DESTROYED without doing anything (the reference's silent no-ops,
 * NRCNetwork.cu:43,66), nrc_b200_reset revives it (Network::init after destroy), nrc_b200_release frees it. */
int nrc_b200_destroy(nrc_b200_handle h);
/* Frees the handle (destroying the network first if it is still alive); the pointer is dead afterwards. */
int nrc_b200_release(nrc_b200_handle h);
/* Network::init on a live object = full reset (new weights, optimizer state zeroed)  Device.cpp:2415-2421 */
int nrc_b200_reset(nrc_b200_handle h, int encoding);

/* Network::infer(inputs_d, outputs_d, numInputs[, stream])  NRCNetwork.cu:64-83.  Asynchronous.
 * Uses the EMA weights (trainer.h:329-335).  `cu_stream` NULL = the stream last set. */
int nrc_b200_infer(nrc_b200_handle h, const float* queries_d, float* radiance_d, uint32_t n, void* cu_stream);
/* Network::train(batchInputs_d, batchTargets_d[, stream], loss_h)  NRCNetwork.cu:41-62.
 * One optimizer step on 16384 samples.  loss_h == NULL: asynchronous.  loss_h != NULL: the loss of this
 * step is written to *loss_h after a stream synchronisation, as in the reference (reduce_sum.h:146-149). */
int nrc_b200_train(nrc_b200_handle h, const float* queries_d, const float* targets_d, void* cu_stream, float* loss_h);

/* Network::setStream  NRCNetwork.cu:85-88 */
int nrc_b200_set_stream(nrc_b200_handle h, void* cu_stream);
/* Network::setHyperParams / getLearningRate  NRCNetwork.cu:90-104 */
int nrc_b200_set_learning_rate(nrc_b200_handle h, float lr);
int nrc_b200_get_learning_rate(nrc_b200_handle h, float* lr);

/* ---- additions without a reference counterpart ------------------------------------------------------ */
/* Arithmetic of the matrix products.  NRC_B200_ACC_FP32 (default): fp32 accumulators in tensor memory, every layer output
 * rounded to fp16 where the reference stores fp16; Hash grid gradients summed in fp32.  NRC_B200_ACC_FP16 = "reference
 * arithmetic": fp16 accumulators, rounded after every 16-wide K block -- what the reference's wmma::accumulator<__half>
 * fragments compute (tcnn fully_fused_mlp.cu:68,150-259) -- in the query kernel (measured bit-identical to the reference's
 * radiance) and in the forward / backward chain of the training step, and Hash grid gradients as half products added with
 * half2 atomics (grid.h:252-255,301-319).  Same speed class; the default is the more accurate arithmetic. */
#define NRC_B200_ACC_FP32 0
#define NRC_B200_ACC_FP16 1
int nrc_b200_set_accumulator(nrc_b200_handle h, int mode);
/* Host-buffer query: chunked H2D / kernel / D2H pipeline inside the call (synchronous).  Page-locked caller
 * buffers are DMA'd directly; pageable ones are staged through an internal pinned ring. */
int nrc_b200_infer_host(nrc_b200_handle h, const float* queries_h, float* radiance_h, uint32_t n);
/* Host-buffer training step (one batch of 16384): H2D on the training stream, then nrc_b200_train.  Asynchronous
 * unless loss_h != NULL; the host buffers must stay valid until the stream has consumed them (pinned memory
 * makes the copies truly asynchronous). */
int nrc_b200_train_host(nrc_b200_handle h, const float* queries_h, const float* targets_h, float* loss_h);
/* Device-side loss of the most recent train() call, without forcing a sync on the training stream. */
int nrc_b200_last_loss_device(nrc_b200_handle h, const float** loss_d);

/* Data-parallel training: a rank computes the fp32 gradient of its shard of a global batch
 * (normalised by 3*global_n) into a device buffer the caller all-reduces (NCCL sum), then every rank applies
 * the identical Adam+EMA step.  grad_d has nrc_b200_n_params() floats. */
int nrc_b200_train_grad(nrc_b200_handle h, const float* queries_d, const float* targets_d, uint32_t n_local,
                        uint32_t n_global, float** grad_d, void* cu_stream);
int nrc_b200_train_apply(nrc_b200_handle h, void* cu_stream, float* loss_h);

/* nrc_b200_train_grad without any allocation or handle-owned scratch.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernel and arithmetic as nrc_b200_train_grad (forward + loss + backward + deterministic
 * weight-gradient reduction, no optimizer step; the handle supplies the weights), but the per-tile gradient partials, the
 * per-tile losses and the Hash grid-gradient accumulator live in a static __device__ array of the library (6.53 MiB, held by
 * every GPU that loads the library until the process exits) and the results go to caller buffers:
 *   grad_out_d  device, nrc_b200_n_params() floats: the fp32 gradient, loss-scaled (x128) like NRC_B200_GRAD_F32
 *   loss_out_d  device, 1 float (may be NULL): the batch loss
 * Limits (NRC_B200_ERR_INVALID with a message naming them otherwise): n_local a multiple of 128 and <= 4096 samples per call
 * (nrc_b200_noalloc_limits reports it), log2_hashmap_size <= 15.  The call performs no cudaMalloc / cudaFree and never
 * synchronises.  All calls on one device share the array, which keeps the last call's contents (what must start at zero is
 * cleared on the call's stream): calls that overlap on the GPU -- other streams, other handles of the same device -- race on
 * it and must be serialised by the caller.  It does not disturb a gradient pending for nrc_b200_train_apply. */
int nrc_b200_train_grad_noalloc(nrc_b200_handle h, const float* queries_d, const float* targets_d, uint32_t n_local,
                                uint32_t n_global, float* grad_out_d, float* loss_out_d, void* cu_stream);
int nrc_b200_noalloc_limits(uint32_t* max_samples, uint64_t* scratch_bytes);

/* Data-parallel training with the gradient all-reduce FUSED into the step kernel (peer stores + flags over NVLink;
 * one process per GPU, CUDA IPC).  Protocol: every rank calls nrc_b200_dp_export, the host all-gathers the
 * NRC_B200_IPC_HANDLE_BYTES-byte handles (any transport), every rank calls nrc_b200_dp_connect with the gathered
 * array, synchronises with its peers once (barrier), and from then on calls nrc_b200_train_dp once per step on
 * every rank.  Replicas must start from identical weights.  World size <= 8 (one NVSwitch domain). */
#define NRC_B200_IPC_HANDLE_BYTES 64
int nrc_b200_dp_export(nrc_b200_handle h, void* ipc_handle_out);
int nrc_b200_dp_connect(nrc_b200_handle h, int rank, int world, const void* ipc_handles);
int nrc_b200_train_dp(nrc_b200_handle h, const float* queries_d, const float* targets_d, uint32_t n_local,
                      uint32_t n_global, void* cu_stream, float* loss_h);
/* The same exchange between `world` handles of ONE process, one per GPU (handles[r] = rank r; peer access instead of CUDA IPC),
 * driven from one host thread or one thread per handle: every entry point makes its handle's device current for the call.
 * On a connected handle nrc_b200_train / _train_batches / _train_shuffled / _update train on the GLOBAL batch of
 * world x 16384 samples (loss and gradient normalised by it) and must be called on every rank; a peer that stays silent for
 * NRC_B200_DP_TIMEOUT_MS (environment, default 30000) makes the next call on the waiting handle return NRC_B200_ERR_CUDA. */
int nrc_b200_dp_connect_local(nrc_b200_handle* handles, int world);

/* ---- NEXT rows: the helper kernels either side of the network in Device::render ----------- */
typedef struct nrc_b200_training_record {       /* nrc::TrainingRecord, neural_radiance_caching.h:57-78 (28 bytes) */
    int propTo;                                  /* next record along the radiance propagation, < 0 = end */
    float localThroughput[3];
    int pixelIndex, tileIndex, propLength;       /* debug fields, carried but unused */
} nrc_b200_training_record;
typedef struct nrc_b200_suffix_end_vertex {      /* nrc::TrainingSuffixEndVertex, neural_radiance_caching.h:81-98 (16 bytes) */
    int startTrainRecord;
    float radianceMask;                          /* 1 = self-training, 0 = unbiased */
    int pixelIndex, tileIndex;
} nrc_b200_suffix_end_vertex;
typedef struct nrc_b200_training_stat { float loss; int num_train_records; } nrc_b200_training_stat;   /* nrc::TrainingStat */

/* accumulate_render_radiance (nrc_helpers.cu:77-129, launched by Device::nrcAccumulateRadiance, Device.cpp:1310-1337).
 * render_mode = nrc::RenderMode (0 Full, 1 NoCache, 2 CacheOnly, 3 CacheFirstVertex, 4 DebugCacheNoThroughputModulation,
 * 5 DebugThroughputOnly); output is a float4-per-pixel device buffer (the reference's GL-interop surface, headless). */
int nrc_b200_accumulate_render_radiance(const float* radiance_d, const float* throughput_d, float* output_rgba_d,
                                        uint32_t n_pixels, int render_mode, uint32_t iteration_index, void* cu_stream);
/* propagate_train_radiance (nrc_helpers.cu:131-224, Device::nrcPropagateRadiance, Device.cpp:1382-1417): per tile,
 * targets[i] += localThroughput[i] * radiance along the propTo chain starting at end_vertices[tile].startTrainRecord. */
int nrc_b200_propagate_train_radiance(const nrc_b200_suffix_end_vertex* end_vertices_d, const float* end_radiance_d,
                                      const nrc_b200_training_record* records_d, float* targets_d, uint32_t n_tiles,
                                      void* cu_stream);
/* Device::nrcTrainRadiance (Device.cpp:1482-1513): n_batches (1..NRC_B200_NUM_BATCHES) consecutive Network::train calls on
 * queries_d + b*16384*15 / targets_d + b*16384*3 as ONE launch (grid barriers between the steps instead of launch
 * boundaries); bit-identical to the loop of nrc_b200_train calls.  loss_h (optional, blocks) = mean of the batch losses. */
int nrc_b200_train_batches(nrc_b200_handle h, const float* queries_d, const float* targets_d, int n_batches,
                           void* cu_stream, float* loss_h);
/* generateRandomPermutationForTrain + permute_train_data (NRCUtil.cu:20-35, nrc_helpers.cu:226-249): dst[d] =
 * src[permutation(d, seed) % min(num_records, 65536)], d < 65536.  The permutation is a keyed 16-bit bijection
 * (nrc_b200_permutation) instead of cuRAND keys + a CUB radix sort. */
uint32_t nrc_b200_permutation(uint32_t i, uint32_t seed);
int nrc_b200_permute_train_data(const float* q_src_d, const float* t_src_d, float* q_dst_d, float* t_dst_d,
                                int num_records, uint32_t seed, void* cu_stream);
/* permute_train_data with the caller's permutation (e.g. the reference's cuRAND + CUB result): dst[d] = src[permutation[d] % n]. */
int nrc_b200_permute_train_data_with(const float* q_src_d, const float* t_src_d, float* q_dst_d, float* t_dst_d,
                                     const int* permutation_d, int num_records, void* cu_stream);
/* Fused: the training step on batch `batch_index` (0..3) of the permuted data without materialising it; identical
 * result to nrc_b200_permute_train_data followed by nrc_b200_train on dst + batch_index*16384. */
int nrc_b200_train_shuffled(nrc_b200_handle h, const float* q_src_d, const float* t_src_d, int num_records, uint32_t seed,
                            uint32_t batch_index, void* cu_stream, float* loss_h);
/* Device::adjustTileSize (Device.cpp:818-828), host side. */
void nrc_b200_adjust_tile_size(int tile_x, int tile_y, int num_train_records, int* new_x, int* new_y);

/* Frame driver: the NRC sequence of Device::render (Device.cpp:2495-2515) in one call, everything asynchronous on the
 * handle's stream; stat != NULL reads the mean loss of the four steps back once (one synchronisation). */
typedef struct nrc_b200_frame {
    const float* queries_d;          /* RadianceQuery[num_pixels + num_tiles]  (bufDynamic.radianceQueriesInference) */
    float* results_d;                /* float3[num_pixels + num_tiles]         (bufDynamic.radianceResultsInference) */
    uint32_t num_pixels, num_tiles;
    const float* throughput_d;       /* float3[num_pixels] lastRenderThroughput, NULL = no accumulation */
    float* output_rgba_d;            /* float4[num_pixels] output buffer */
    int render_mode;                 /* nrc::RenderMode */
    uint32_t iteration_index;
    const nrc_b200_suffix_end_vertex* end_vertices_d;   /* [num_tiles] */
    const nrc_b200_training_record* records_d;          /* [65536] */
    const float* train_queries_d;    /* RadianceQuery[65536] (unshuffled, buffer 0) */
    float* train_targets_d;          /* float3[65536] (unshuffled; propagation adds into it) */
    int num_train_records;
    uint32_t shuffle_seed;           /* e.g. the frame index */
} nrc_b200_frame;
int nrc_b200_update(nrc_b200_handle h, const nrc_b200_frame* frame, void* cu_stream, nrc_b200_training_stat* stat);

/* Parity-harness access to the parameters (Trainer::params() / set_params(), trainer.h:226-269). */
typedef enum nrc_b200_param_kind {
    NRC_B200_PARAMS_MASTER_F32 = 0,  /* fp32 master weights  */
    NRC_B200_PARAMS_TRAIN_F16 = 1,   /* fp16 training weights */
    NRC_B200_PARAMS_EMA_F16 = 2,     /* fp16 EMA = inference weights */
    NRC_B200_PARAMS_GRAD_F16 = 3,    /* fp16 gradient of the last step (loss-scaled by 128) */
    NRC_B200_ADAM_M1_F32 = 4,
    NRC_B200_ADAM_M2_F32 = 5,
    NRC_B200_GRAD_F32 = 6            /* fp32 gradient vector of the last train_grad (before the fp16 rounding) */
} nrc_b200_param_kind;
int nrc_b200_n_params(nrc_b200_handle h, uint32_t* n_params, uint32_t* n_mlp_params);
int nrc_b200_get_params(nrc_b200_handle h, int kind, void* dst_host);   /* synchronises the stream */
int nrc_b200_set_params(nrc_b200_handle h, int kind, const void* src_host);
/* Loads fp32 master weights, derives the fp16 training copy, zeroes EMA/Adam state and the step counter. */
int nrc_b200_load_master(nrc_b200_handle h, const float* master_host);

/* Debug aid: per-tile clock64 phase stamps of the training step kernel ([128][8] int64; see csrc/nrc_b200.cu). */
int nrc_b200_debug_train_timeline(nrc_b200_handle h, int enable, long long* dst_host);

/* Measurement aid: GB/s the SMs read from an L2-resident buffer of `bytes` (16-byte L1-bypassing loads, best of 5 x `reps`
 * passes) on the current device -- the measured denominator of the hash-grid gather roofline in bench.py. */
int nrc_b200_debug_l2_read_bandwidth(uint64_t bytes, int reps, float* gbs);

const char* nrc_b200_last_error(void);
/* Number of kernel launches issued through this handle so far (bench.py's gpu_launches). */
int nrc_b200_launch_count(nrc_b200_handle h, uint64_t* count);

#ifdef __cplusplus
}
#endif
#endif
