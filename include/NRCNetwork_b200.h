// NRCNetwork_b200.h -- header-only re-creation of the reference's cache interface `class nrc::Network`
// (/root/reference/nrc/inc/NRCNetwork.h:9-76, behaviour of nrc/src/NRCNetwork.cu:22-128) on top of the
// C ABI in nrc_b200.h.  Drop-in: replace `#include "inc/NRCNetwork.h"` by this header (or copy it over
// inc/NRCNetwork.h), drop nrc/src/NRCNetwork.cu and tiny-cuda-nn from the build, link libnrc_b200.so.
// nrc/src/Device.cpp compiles unchanged: same class name, method names, signatures, default arguments,
// PODs, exception types and the "silent no-op after destroy()" behaviour.
#pragma once
#include "nrc_b200.h"

#include <cstdint>
#include <iostream>
#include <limits>
#include <stdexcept>
#include <string>

#if defined(__has_include)
#if __has_include("shaders/neural_radiance_caching.h") && !defined(NRC_B200_STANDALONE)
#include "shaders/config.h"
#include "shaders/neural_radiance_caching.h"   // nrc::InputEncoding, BATCH_SIZE, NN_INPUT_DIMS, TRAIN_LR ...
#define NRC_B200_HAVE_REFERENCE_TYPES 1
#endif
#endif

typedef struct CUstream_st* CUstream;            // identical to <cuda.h>'s typedef (benign redeclaration)

namespace nrc {
#ifndef NRC_B200_HAVE_REFERENCE_TYPES
    // Stand-alone builds (tests, other hosts): the few types of shaders/neural_radiance_caching.h:24-55 we need.
    enum class InputEncoding : int { Frequency = 0, Hash };
    constexpr int NUM_BATCHES = NRC_B200_NUM_BATCHES;
    constexpr int BATCH_SIZE = NRC_B200_BATCH_SIZE;
    constexpr int NN_INPUT_DIMS = NRC_B200_INPUT_DIMS;
    constexpr int NN_OUTPUT_DIMS = NRC_B200_OUTPUT_DIMS;
    constexpr float TRAIN_LR(InputEncoding e) {
        return e == InputEncoding::Frequency ? 1e-3f : e == InputEncoding::Hash ? 1e-2f : 1e-4f;
    }
#endif

    struct HyperParams {            // NRCNetwork.h:10-13
        float learningRate;
    };

    struct TrainingStat {           // NRCNetwork.h:15-18
        float loss{std::numeric_limits<float>::quiet_NaN()};
        int numTrainRecords{0};
    };

    class Network {
    public:
        Network() = default;

        ~Network()                  // NRCNetwork.cu:29-33: the destructor only warns (device memory is NOT freed here)
        {
            if (!m_destroyed)
                std::cerr << "WARNING: NRC Network must be explicitly destroy() in the context it was created in!\n";
            else if (m_handle)
                nrc_b200_release(m_handle);     // host-side tombstone of a destroyed network: no CUDA call involved
        }

        Network(const Network&) = delete;
        Network& operator=(const Network&) = delete;

        template<bool Verbose = false>
        void init(CUstream stream, nrc::InputEncoding encoding)   // NRCNetwork.h:26-31, NRCNetwork.cu:106-112
        {
            m_stream = stream;
            if (m_handle) {         // init on a live or destroyed object = full re-create (Device.cpp:2415-2421)
                check(nrc_b200_reset(m_handle, static_cast<int>(encoding)));
                check(nrc_b200_set_stream(m_handle, stream));
            } else {
                check(nrc_b200_create(static_cast<int>(encoding), stream, &m_handle));
            }
            m_encoding = encoding;
            m_destroyed = false;
            if constexpr (Verbose) printConfig_();
        }

        void destroy()              // NRCNetwork.cu:35-39
        {
            // the C handle stays behind as a tombstone: calls after destroy() are no-ops inside the library as well
            if (m_handle) nrc_b200_destroy(m_handle);
            m_destroyed = true;
        }

        // Perform a single training step (NRCNetwork.cu:41-62).  loss_h != nullptr blocks on the stream.
        void train(float* batchInputs_d, float* batchTargets_d, float* loss_h = nullptr)
        {
            if (m_destroyed || !m_handle) return;
            check(nrc_b200_train(m_handle, batchInputs_d, batchTargets_d, m_stream, loss_h));
        }

        void train(float* batchInputs_d, float* batchTargets_d, CUstream stream, float* loss_h = nullptr)
        {
            setStream(stream);
            train(batchInputs_d, batchTargets_d, loss_h);
        }

        // Perform inference on the input (NRCNetwork.cu:64-83).  Rows [numInputs, round_up(numInputs,256)) of the
        // caller's buffers are NOT touched (the reference reads and writes them).
        void infer(float* inputs_d, float* outputs_d, uint32_t numInputs)
        {
            if (m_destroyed || !m_handle) return;
            check(nrc_b200_infer(m_handle, inputs_d, outputs_d, numInputs, m_stream));
        }

        void infer(float* inputs_d, float* outputs_d, uint32_t numInputs, CUstream stream)
        {
            setStream(stream);
            infer(inputs_d, outputs_d, numInputs);
        }

        void setStream(CUstream stream)                           // NRCNetwork.cu:85-88
        {
            m_stream = stream;
            if (m_handle) check(nrc_b200_set_stream(m_handle, stream));
        }

        void setHyperParams(const HyperParams& hp)                // NRCNetwork.cu:90-94
        {
            if (m_handle) check(nrc_b200_set_learning_rate(m_handle, hp.learningRate));
        }

        void setConfig(InputEncoding encoding) { m_encoding = encoding; }   // takes effect at the next init()

        float getLearningRate() const                             // NRCNetwork.cu:101-104
        {
            float lr = TRAIN_LR(m_encoding);
            if (m_handle) check(nrc_b200_get_learning_rate(m_handle, &lr));
            return lr;
        }

        nrc_b200_handle handle() const { return m_handle; }       // escape hatch to the extended C ABI

    private:
        nrc_b200_handle m_handle{nullptr};
        CUstream m_stream{};
        InputEncoding m_encoding{InputEncoding::Frequency};
        bool m_destroyed{false};

        static void check(int rc)
        {
            if (rc == NRC_B200_OK || rc == NRC_B200_ERR_DESTROYED) return;
            const std::string msg = nrc_b200_last_error();
            if (rc == NRC_B200_ERR_INVALID) throw std::invalid_argument(msg);   // NRCNetworkConfigs.h:129-131
            throw std::runtime_error(msg);                                       // CUDA_CHECK_THROW / CHECK_THROW
        }

        void printConfig_() const                                 // NRCNetwork.cu:122-127
        {
            uint32_t n = 0, n_mlp = 0;
            if (m_handle) nrc_b200_n_params(m_handle, &n, &n_mlp);
            std::cout << "\n----------------------- NETWORK CONFIG -----------------------\n"
                      << "encoding: " << (m_encoding == InputEncoding::Frequency
                                              ? "TriangleWave(12)x3 + OneBlob(4)x6 + Identity x6 (-> 80)"
                                              : "HashGrid(16 levels x 2, 2^15) + OneBlob(4)x6 + Identity x6 (-> 64)")
                      << "\nnetwork: FullyFusedMLP 5x64 ReLU, ReLU output, 3 (padded 16) outputs"
                      << "\nloss: RelativeL2Luminance   optimizer: EMA(0.99) o Adam(lr " << getLearningRate()
                      << ", l2_reg 1e-6)\nparameters: " << n << " (" << n_mlp << " MLP)"
                      << "\n--------------------------------------------------------------\n\n";
        }
    };
}
