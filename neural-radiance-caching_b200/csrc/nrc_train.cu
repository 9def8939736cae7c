// placeholder until the fused training kernels land (next commit)
#include "nrc_kernels.h"
namespace nrcb {
cudaError_t launch_train_fwd_bwd(int, const float*, const float*, uint32_t, uint32_t, const TrainBuffers&, const GridMeta&, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t launch_reduce_grads(const TrainBuffers&, uint32_t, uint32_t, uint32_t, float*, float*, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t launch_adam_ema(const OptimState&, const float*, float*, uint32_t, cudaStream_t) { return cudaErrorNotSupported; }
}
