// Empty stand-in for <optix.h>: nrc/shaders/nrc_helpers.cu includes it but its three plain CUDA kernels
// (accumulate_render_radiance, propagate_train_radiance, permute_train_data) use nothing from OptiX.
// Used ONLY by oracle/ref/Makefile.ref to compile those kernels, unmodified, into oracle/_ref.
#pragma once
