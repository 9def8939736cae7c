// Stub for the reference's nrc/shaders/config.h, used ONLY to compile the reference's
// nrc/src/NRCNetwork.cu + tiny-cuda-nn without OptiX/MDL/GL (oracle/ref/Makefile.ref).
// It restates the single switch NRCNetwork*.h depend on (config.h:113).
#pragma once
#include <cuda_runtime.h>
#define USE_COMPACT_RADIANCE_QUERY 1
