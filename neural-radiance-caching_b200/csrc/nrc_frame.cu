// The NRC helper kernels that sit either side of the network in Device::render (SURVEY.md 8f "NEXT" rows),
// re-implemented as plain CUDA for sm_100a:
//   accumulate_render_radiance  nrc/shaders/nrc_helpers.cu:77-129   (output += throughput * radiance / (iteration + 1))
//   propagate_train_radiance    nrc/shaders/nrc_helpers.cu:131-224  (walk the TrainingRecord chains of every tile)
//   permute_train_data          nrc/shaders/nrc_helpers.cu:226-249  (+ the permutation of nrc/src/NRCUtil.cu:20-35)
// The reference compiles them into its OptiX module and reads launch state from the `sysData` constant; here every
// input is an explicit argument and the output surface is a headless float4 device buffer.
#include <cuda_runtime.h>
#include "nrc_kernels.h"
#include "nrc_frame.cuh"

namespace nrcb {

// One thread per pixel; float3 arrays are read as 3 scalars (12-byte records are not vector-aligned), the float4
// output is one 16-byte access.  HBM-bound: 12 + 12 + 16 B read, 16 B written per pixel.
__global__ void __launch_bounds__(256)
nrc_accumulate_kernel(const float* __restrict__ radiance, const float* __restrict__ throughput, float4* __restrict__ out,
                      uint32_t n_pixels, int mode, float acc_weight) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const float r0 = radiance[3 * i], r1 = radiance[3 * i + 1], r2 = radiance[3 * i + 2];
    const float t0 = throughput[3 * i], t1 = throughput[3 * i + 1], t2 = throughput[3 * i + 2];
    switch (mode) {
        case 0: {   // RenderMode::Full
            const float4 d = out[i];
            out[i] = make_float4(fmaf(t0 * r0, acc_weight, d.x), fmaf(t1 * r1, acc_weight, d.y), fmaf(t2 * r2, acc_weight, d.z), 1.0f);
            break;
        }
        case 2: out[i] = make_float4(r0 * t0, r1 * t1, r2 * t2, 1.0f); break;   // CacheOnly
        case 4: out[i] = make_float4(r0, r1, r2, 1.0f); break;                   // DebugCacheNoThroughputModulation
        case 5: out[i] = make_float4(t0, t1, t2, 1.0f); break;                   // DebugThroughputOnly
        default: break;                                                          // NoCache / CacheFirstVertex: nothing
    }
}

// One thread per tile: target[i] += localThroughput[i] * radiance carried from the previous vertex, along propTo links
// (chains are short: bounded by the path length).  Records are 28 bytes, end vertices 16 bytes.
__global__ void __launch_bounds__(128)
nrc_propagate_kernel(const int4* __restrict__ end_vertices, const float* __restrict__ end_radiance,
                     const int* __restrict__ records, float* __restrict__ targets, uint32_t n_tiles, uint32_t n_records_cap) {
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= n_tiles) return;
    const int4 ev = end_vertices[tile];                  // {startTrainRecord, radianceMask, pixelIndex, tileIndex}
    const float mask = __int_as_float(ev.y);
    float l0 = end_radiance[3 * tile] * mask, l1 = end_radiance[3 * tile + 1] * mask, l2 = end_radiance[3 * tile + 2] * mask;
    int i = ev.x;
    uint32_t guard = 0;
    while (i >= 0 && (uint32_t)i < n_records_cap && guard++ < 4096) {
        const int* rec = records + (size_t)i * 7;        // {propTo, localThroughput.xyz, pixelIndex, tileIndex, propLength}
        float* t = targets + (size_t)i * 3;
        l0 = fmaf(__int_as_float(rec[1]), l0, t[0]);
        l1 = fmaf(__int_as_float(rec[2]), l1, t[1]);
        l2 = fmaf(__int_as_float(rec[3]), l2, t[2]);
        t[0] = l0; t[1] = l1; t[2] = l2;
        i = rec[0];
    }
}

// dst[d] = src[permutation(d) % num_records] for the 65536 training slots (queries: 15 floats, targets: 3 floats).
// One thread per float of the destination: fully coalesced stores, 60-byte gathers.
__global__ void __launch_bounds__(256)
nrc_permute_kernel(const float* __restrict__ q_src, const float* __restrict__ t_src, float* __restrict__ q_dst,
                   float* __restrict__ t_dst, uint32_t num_records, uint32_t seed) {
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;   // 0 .. 65536*18
    if (e >= 65536u * 18u) return;
    const uint32_t d = e / 18u, c = e - d * 18u;
    const uint32_t s = permute16(d, seed) % num_records;
    if (c < 15u) q_dst[d * 15u + c] = __ldg(q_src + (size_t)s * 15u + c);
    else t_dst[d * 3u + (c - 15u)] = __ldg(t_src + (size_t)s * 3u + (c - 15u));
}

// The same gather with a caller-supplied permutation (the reference's cuRAND keys + CUB radix sort, NRCUtil.cu:20-35):
// permute_train_data of nrc_helpers.cu:226-249 argument for argument.
__global__ void __launch_bounds__(256)
nrc_permute_with_kernel(const float* __restrict__ q_src, const float* __restrict__ t_src, float* __restrict__ q_dst,
                        float* __restrict__ t_dst, const int* __restrict__ permutation, uint32_t num_records) {
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;   // 0 .. 65536*18
    if (e >= 65536u * 18u) return;
    const uint32_t d = e / 18u, c = e - d * 18u;
    const int s = __ldg(permutation + d) % (int)num_records;   // signed modulo, as the reference computes it
    if (s < 0) return;
    if (c < 15u) q_dst[d * 15u + c] = __ldg(q_src + (size_t)s * 15u + c);
    else t_dst[d * 3u + (c - 15u)] = __ldg(t_src + (size_t)s * 3u + (c - 15u));
}

// Measurement aid (bench.py's L2 roofline denominator): every thread sums 16-byte L1-bypassing loads over an L2-resident buffer.
__global__ void __launch_bounds__(256)
nrc_l2_read_probe_kernel(const float4* __restrict__ buf, size_t n_vec, int reps, float* sink) {
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += 4 * stride) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = i + u * stride < n_vec ? __ldcg(buf + i + u * stride) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
            for (int u = 0; u < 4; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
        }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) *sink = acc.x;     // never true: keeps the loads alive
}
cudaError_t l2_read_probe(size_t bytes, int reps, float* gbs) {
    float4* buf = nullptr; float* sink = nullptr;
    cudaError_t e = cudaMalloc(&buf, bytes);
    if (e != cudaSuccess) return e;
    cudaMalloc(&sink, 4);
    cudaMemset(buf, 0, bytes);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    nrc_l2_read_probe_kernel<<<sms * 8, 256>>>(buf, bytes / 16, 2, sink);       // warm-up: brings the buffer into L2
    float best = 1e30f;
    for (int it = 0; it < 5; ++it) {
        cudaEventRecord(a);
        nrc_l2_read_probe_kernel<<<sms * 8, 256>>>(buf, bytes / 16, reps, sink);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    e = cudaGetLastError();
    *gbs = (float)((double)bytes * reps / (best * 1e-3) / 1e9);
    cudaEventDestroy(a); cudaEventDestroy(b);
    cudaFree(buf); cudaFree(sink);
    return e;
}

cudaError_t launch_accumulate(const float* radiance, const float* throughput, float* out4, uint32_t n_pixels, int mode,
                              uint32_t iteration_index, cudaStream_t stream) {
    if (n_pixels == 0) return cudaSuccess;
    nrc_accumulate_kernel<<<(n_pixels + 255) / 256, 256, 0, stream>>>(radiance, throughput, reinterpret_cast<float4*>(out4), n_pixels,
                                                                      mode, 1.0f / (float)(iteration_index + 1));
    return cudaGetLastError();
}
cudaError_t launch_propagate(const void* end_vertices, const float* end_radiance, const void* records, float* targets,
                             uint32_t n_tiles, cudaStream_t stream) {
    if (n_tiles == 0) return cudaSuccess;
    nrc_propagate_kernel<<<(n_tiles + 127) / 128, 128, 0, stream>>>(reinterpret_cast<const int4*>(end_vertices), end_radiance,
                                                                    reinterpret_cast<const int*>(records), targets, n_tiles, 65536u);
    return cudaGetLastError();
}
cudaError_t launch_permute(const float* q_src, const float* t_src, float* q_dst, float* t_dst, uint32_t num_records,
                           uint32_t seed, cudaStream_t stream) {
    nrc_permute_kernel<<<(65536u * 18u + 255) / 256, 256, 0, stream>>>(q_src, t_src, q_dst, t_dst, num_records, seed);
    return cudaGetLastError();
}

cudaError_t launch_permute_with(const float* q_src, const float* t_src, float* q_dst, float* t_dst, const int* permutation,
                                uint32_t num_records, cudaStream_t stream) {
    nrc_permute_with_kernel<<<(65536u * 18u + 255) / 256, 256, 0, stream>>>(q_src, t_src, q_dst, t_dst, permutation, num_records);
    return cudaGetLastError();
}

}  // namespace nrcb
