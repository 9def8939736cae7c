// Force-included (-include) in front of the reference's nrc/shaders/nrc_helpers.cu by oracle/ref/Makefile.ref.
// The reference's system_data.h / per_ray_data.h pull in the MDL SDK and OptiX device headers, which are not on this box;
// the helper kernels touch five scalars of `sysData` (nrc_helpers.cu:42-45,84-104,144-149,237) and the inline function
// endTrainSuffixUnbiased touches two fields of PerRayData (neural_radiance_caching.h:207-221).  This header pre-defines the
// include guards of those two headers and declares exactly those members, so that the kernels' SOURCE is compiled as it is.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#define SYSTEM_DATA_H
#define PER_RAY_DATA_H
namespace nrc { struct ControlBlock; }
struct SystemDataPerFrame {            // system_data.h:56-75 (members the helper kernels read)
    int2 tileSize;
    int2 numTiles;
    int iterationIndex;
};
struct SystemData {                    // system_data.h:80-138 (members the helper kernels read)
    nrc::ControlBlock* nrcCB;
    CUdeviceptr outputBuffer;
    int2 resolution;
    SystemDataPerFrame pf;
};
struct PerRayData {                    // per_ray_data.h (members neural_radiance_caching.h:207-221 reads)
    int tileIndex;
    int lastTrainRecordIndex;
    int pixelIndex;
};
