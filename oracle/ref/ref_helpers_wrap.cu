// C wrapper around the reference's UNMODIFIED frame helper kernels -- accumulate_render_radiance,
// propagate_train_radiance, permute_train_data (/root/reference/nrc/shaders/nrc_helpers.cu:77-249) -- so that tests can
// pin the NEXT rows by the reference's own outputs.  The kernels' source is compiled where it lies
// (NRC_HELPERS_CU is its path, given by oracle/ref/Makefile.ref); this TU only fills `sysData` and launches them the way
// Device::nrcAccumulateRadiance / nrcPropagateRadiance / nrcShuffleTrainingData do (Device.cpp:1310-1337,1382-1417,1440-1471).
// Test infrastructure only.
// nrc_helpers.cu only DECLARES the launch-state constant (`extern "C" __constant__ SystemData sysData;`: in the reference the
// definition comes with the module the renderer loads); this TU defines it.
extern "C" { __constant__ SystemData sysData; }
#include NRC_HELPERS_CU
#include <cstdint>
#include <string>

static thread_local std::string g_herr;
static int herr(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    g_herr = std::string(what) + ": " + cudaGetErrorString(e);
    return -1;
}
#define HCK(x) do { int r_ = herr((x), #x); if (r_) return r_; } while (0)

static int set_sys(const SystemData& s) { return herr(cudaMemcpyToSymbol(sysData, &s, sizeof(s)), "cudaMemcpyToSymbol(sysData)"); }

extern "C" {
const char* nrc_refh_last_error() { return g_herr.c_str(); }

// output: float4[width*height]; radiance / throughput: float3[width*height]
int nrc_refh_accumulate(float* radiance_d, float* throughput_d, float* output_d, int width, int height, int mode, int iteration) {
    SystemData s{};
    s.outputBuffer = (CUdeviceptr)output_d;
    s.resolution = make_int2(width, height);
    s.pf.iterationIndex = iteration;
    if (set_sys(s)) return -1;
    const dim3 block(16, 16), grid((width + 15) / 16, (height + 15) / 16);
    accumulate_render_radiance<<<grid, block>>>((float3*)radiance_d, (float3*)throughput_d, (nrc::RenderMode)mode);
    HCK(cudaGetLastError());
    HCK(cudaDeviceSynchronize());
    return 0;
}
int nrc_refh_propagate(void* end_vertices_d, float* end_radiance_d, void* records_d, float* targets_d, int tiles_x, int tiles_y) {
    SystemData s{};
    s.pf.numTiles = make_int2(tiles_x, tiles_y);
    if (set_sys(s)) return -1;
    const dim3 block(16, 16), grid((tiles_x + 15) / 16, (tiles_y + 15) / 16);
    propagate_train_radiance<<<grid, block>>>((nrc::TrainingSuffixEndVertex*)end_vertices_d, (float3*)end_radiance_d,
                                              (nrc::TrainingRecord*)records_d, (float3*)targets_d);
    HCK(cudaGetLastError());
    HCK(cudaDeviceSynchronize());
    return 0;
}
int nrc_refh_permute(float* q_src_d, float* q_dst_d, float* t_src_d, float* t_dst_d, int* permutation_d, int num_records) {
    nrc::ControlBlock cb{};
    cb.numTrainingRecords = num_records;
    nrc::ControlBlock* cb_d = nullptr;
    HCK(cudaMalloc(&cb_d, sizeof(cb)));
    HCK(cudaMemcpy(cb_d, &cb, sizeof(cb), cudaMemcpyHostToDevice));
    SystemData s{};
    s.nrcCB = cb_d;
    if (set_sys(s)) { cudaFree(cb_d); return -1; }
    nrc::DoubleBuffer<nrc::RadianceQuery> qb;
    qb.data[0] = (nrc::RadianceQuery*)q_src_d; qb.data[1] = (nrc::RadianceQuery*)q_dst_d;
    nrc::DoubleBuffer<float3> tb;
    tb.data[0] = (float3*)t_src_d; tb.data[1] = (float3*)t_dst_d;
    permute_train_data<<<nrc::NUM_TRAINING_RECORDS_PER_FRAME / 256, 256>>>(qb, tb, permutation_d);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaFree(cb_d);
    return herr(e, "permute_train_data");
}
}
