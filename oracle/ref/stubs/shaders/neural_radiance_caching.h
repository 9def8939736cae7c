// Stub for the reference's nrc/shaders/neural_radiance_caching.h (which pulls in OptiX and MDL
// headers that are not on this box).  Restates only the constants nrc::Network needs
// (neural_radiance_caching.h:24-54); used ONLY by oracle/ref/Makefile.ref.
#pragma once
namespace nrc {
    enum class InputEncoding : int { Frequency = 0, Hash };
    constexpr int NUM_BATCHES = 4;
    constexpr int NUM_TRAINING_RECORDS_PER_FRAME = 65536;
    constexpr int BATCH_SIZE = NUM_TRAINING_RECORDS_PER_FRAME / NUM_BATCHES;
    constexpr int NN_INPUT_DIMS = 3 + 2 + 2 + 2 + 3 + 3;
    constexpr int NN_OUTPUT_DIMS = 3;
    constexpr float TRAIN_LR(InputEncoding encoding)
    {
        switch (encoding) {
            case InputEncoding::Frequency: return 1e-3f;
            case InputEncoding::Hash: return 1e-2f;
            default: return 1e-4f;
        }
    }
}
