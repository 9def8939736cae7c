// Pieces of the per-frame NRC sequence next to the network, shared by host and device code.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define NRCB_HD __host__ __device__ __forceinline__
#else
#define NRCB_HD inline
#endif

namespace nrcb {

// Per-frame pseudo-random permutation of the 65536 training-record slots.
// The reference draws 65536 random keys with cuRAND and radix-sorts (key, index) pairs with CUB
// (nrc/src/NRCUtil.cu:20-35) only to decorrelate neighbouring records; NUM_TRAINING_RECORDS_PER_FRAME is exactly 2^16, so
// a keyed bijection on 16 bits gives a permutation in O(1) per element: no keys, no sort, no buffers, and the gather
// can be fused into the training kernel's loader.  4-round Feistel network on two 8-bit halves.
NRCB_HD uint32_t feistel_round(uint32_t x, uint32_t key) {
    x = (x + key) * 0x9E3779B1u;
    x ^= x >> 15;
    x *= 0x85EBCA77u;
    x ^= x >> 13;
    return x & 0xFFu;
}
NRCB_HD uint32_t permute16(uint32_t i, uint32_t seed) {
    uint32_t l = (i >> 8) & 0xFFu, r = i & 0xFFu;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (uint32_t k = 0; k < 4; ++k) {
        const uint32_t t = l ^ feistel_round(r, seed * 4u + k + 0x632BE5ABu * (k + 1u));
        l = r;
        r = t;
    }
    return (l << 8) | r;
}

}  // namespace nrcb
