// C wrapper around the UNMODIFIED reference cache (nrc::Network, /root/reference/nrc/src/NRCNetwork.cu)
// so tests and bench.py can drive it through ctypes on the GPU box.  Test infrastructure only.
// nrc::Network hides its tcnn model behind a private PIMPL; the parity harness needs the weights,
// so this TU re-opens the class (private -> public) and repeats the Impl definition verbatim in
// shape (NRCNetwork.cu:15-20) to reach model.trainer.
#include <limits>
#define private public
#include "inc/NRCNetwork.h"
#undef private
#include <tiny-cuda-nn/common.h>
#include <tiny-cuda-nn/config.h>
#include <cstring>
#include <string>

namespace nrc {
struct Network::Impl {
    tcnn::TrainableModel model{};
    nlohmann::json config{};
};
}

static thread_local std::string g_err;
#define GUARD(body)                          \
    try { body; return 0; }                  \
    catch (const std::exception& e) { g_err = e.what(); return -1; }

extern "C" {
const char* nrc_ref_last_error() { return g_err.c_str(); }
int nrc_ref_create(int encoding, void* stream, void** out) {
    GUARD({
        auto* n = new nrc::Network();
        n->init<false>((CUstream)stream, (nrc::InputEncoding)encoding);
        *out = n;
    })
}
int nrc_ref_destroy(void* h) { GUARD({ auto* n = (nrc::Network*)h; n->destroy(); delete n; }) }
int nrc_ref_infer(void* h, float* in_d, float* out_d, uint32_t n, void* stream) {
    GUARD(((nrc::Network*)h)->infer(in_d, out_d, n, (CUstream)stream))
}
int nrc_ref_train(void* h, float* in_d, float* tgt_d, void* stream, float* loss_h) {
    GUARD(((nrc::Network*)h)->train(in_d, tgt_d, (CUstream)stream, loss_h))
}
int nrc_ref_set_lr(void* h, float lr) { GUARD(((nrc::Network*)h)->setHyperParams(nrc::HyperParams{lr})) }
int nrc_ref_n_params(void* h) { return (int)((nrc::Network*)h)->pImpl->model.trainer->n_params(); }
// which: 0 = fp32 master, 1 = fp16 training weights, 2 = fp16 inference (EMA) weights,
//        3 = fp16 gradients of the last step (loss-scaled by 128); host buffers
int nrc_ref_get_params(void* h, int which, void* dst_h) {
    GUARD({
        auto& t = ((nrc::Network*)h)->pImpl->model.trainer;
        size_t n = t->n_params();
        const void* src = which == 0 ? (const void*)t->params_full_precision()
                        : which == 1 ? (const void*)t->params()
                        : which == 2 ? (const void*)t->params_inference() : (const void*)t->param_gradients();
        CUDA_CHECK_THROW(cudaDeviceSynchronize());
        CUDA_CHECK_THROW(cudaMemcpy(dst_h, src, n * (which == 0 ? 4 : 2), cudaMemcpyDeviceToHost));
    })
}
// sets fp32 master + fp16 training weights (Trainer::set_params_full_precision, which also writes the fp16 copy through
// m_params_inference, i.e. the EMA buffer the network infers with, trainer.h:242-254); the Adam moments are left as they are
// (fresh model => zeros).  The EMA filter itself starts from debias_old = 0 at step 1, so its first output does not depend on
// what the buffer held.
int nrc_ref_set_master(void* h, const float* src_h) {
    GUARD({
        auto& t = ((nrc::Network*)h)->pImpl->model.trainer;
        t->set_params_full_precision(src_h, t->n_params(), false);
        CUDA_CHECK_THROW(cudaDeviceSynchronize());
    })
}
}
