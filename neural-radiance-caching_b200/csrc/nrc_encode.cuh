// Input encodings of the NRC network, evaluated per query in registers (never written to HBM).
// Math follows the reference's tiny-cuda-nn kernels ("tcnn/" = 3rdparty/tiny-cuda-nn/include/tiny-cuda-nn):
//   TriangleWave  tcnn/encodings/triangle_wave.h:61-76
//   OneBlob       tcnn/encodings/oneblob.h:47-67 (AoS, Frequency config) / :99-127 (SoA, Hash config),
//                 quartic_cdf tcnn/common_device.h:915-920
//   Identity+pad  tcnn/encodings/identity.h:59-67 (padding lanes are 1.0)
//   HashGrid      tcnn/encodings/grid.h:94-169, tcnn/common_device.h:651-707,856-868
// Feature order = CompositeEncoding row order (composite.h:214-280):
//   Frequency: [0,36) tri-wave d*12+f | [36,60) one-blob | [60,66) albedos | [66,80) 1.0
//   Hash     : [0,32) grid level*2+f  | [32,56) one-blob | [56,62) albedos | [62,64) 1.0
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include "sm100_prims.cuh"

namespace nrcb {

// Layout of the multiresolution hash grid (grid.h:692-723), filled by the host (nrc_b200.cu).
struct GridMeta {
    uint32_t offset[17];   // cumulative entry offsets (entries of 2 fp16 features)
    uint32_t stride_y[16]; // index = x + y*stride_y + z*stride_z (mod 2^32) on the levels that are not hashed: the strides the
    uint32_t stride_z[16]; // reference's grid_index loop ends up with, INCLUDING its uint32 wrap (common_device.h:691-707)
    float scale[16];
    uint32_t dense_mask;   // bit l set: level l is NOT hashed (res^3 <= table size, or the stride wrapped to 0)
};

// NRC_ENC_VARIANT selects how the Frequency / one-blob arithmetic is issued (same mathematics):
//   0  scalar fp32, FRND for floor / rint (the first version of this round)
//   1  tri-wave folded to 1 - 4|v - rint(v)|, rint through the 1.5*2^23 magic add (no XU pipe), scalar
//   2  = 1 with two features per instruction: Blackwell packed fp32 (FFMA2 / FADD2 / FMUL2, one issue slot for two lanes)
//   3  = 2 plus the one-blob CDF polynomials two bin edges at a time
#ifndef NRC_ENC_VARIANT
#define NRC_ENC_VARIANT 3
#endif

// Packed fp32 pairs live in one 64-bit register (sm_100+: fma/add/mul .f32x2, each lane IEEE round-to-nearest).
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_make(float lo, float hi) {
    f2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void f2_split(f2_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) {
    f2_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) {
    f2_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) {
    f2_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
#define NRC_RINT_MAGIC 12582912.0f   // 1.5 * 2^23: (v + M) - M == rint(v) for |v| < 2^22

// Wrapped quartic-kernel CDF sum  c(d) + c(d-1) + c(d+1)  with  c(t) = clamp(P(4t) + 0.5, 0, 1).
// c saturates for |t| >= 1/4, so at most one of the three terms is unsaturated: with k = rint(d), r = d - k
// the sum is  clamp(k+1, 0, 3) + (|k| <= 1 ? c(r) : 0).  r is exact, so the live term matches the
// reference's bit for bit; the saturated terms contribute exact integers.
__device__ __forceinline__ float blob_cdf3(float d) {
    const float k = rintf(d);
    const float r = d - k;
    // c(r) = sat(15/16 u (1 - 2/3 u^2 + 1/5 u^4) + 1/2), u = 4r.  Powers of two are folded into the constants
    // (u^2 = 16 r^2, u^4 = 256 r^4, 15/16 u = 3.75 r): every product keeps the reference's rounding exactly.
    const float q2 = r * r;
    const float q4 = q2 * q2;
    const float inner = fmaf(q4, 256.0f * 0.2f, fmaf(q2, 16.0f * -(2.0f / 3.0f), 1.0f));
    const float c = __saturatef(fmaf(3.75f * r, inner, 0.5f));
    // clamp(k+1, 0, 3) + (|k| <= 1 ? c : 0)  ==  clamp((k+1) + c, 0, 3):  for |k| >= 2 the sum is already outside [0, 3].
    return fminf(fmaxf((k + 1.0f) + c, 0.0f), 3.0f);
}

// 4 bins of one dimension. WRAP_LAST = true: Frequency/AoS path (last bin's right edge = left edge of bin 0
// plus 1, oneblob.h:60-64); false: Hash/SoA path (right edge evaluated at 1.0, oneblob.h:117-121).
template <bool WRAP_LAST>
__device__ __forceinline__ void oneblob4(float x, float (&y)[4]) {
    const float c0 = blob_cdf3(0.0f - x);
    const float c1 = blob_cdf3(0.25f - x);
    const float c2 = blob_cdf3(0.5f - x);
    const float c3 = blob_cdf3(0.75f - x);
    const float c4 = WRAP_LAST ? c0 + 1.0f : blob_cdf3(1.0f - x);
    y[0] = c1 - c0; y[1] = c2 - c1; y[2] = c3 - c2; y[3] = c4 - c3;
}

// Packed form of blob_cdf3 for two dimensions at once (lanes = dims), bin edge b broadcast: the same IEEE operations as
// blob_cdf3 lane by lane (b - x as fma(x, -1, b): one rounding; rint through the magic add: exact for |b - x| < 2^22),
// so the results are bit-identical to the scalar form.  The saturating fma and the clamp have no packed form: scalar.
__device__ __forceinline__ void blob_cdf3_x2(f2_t x2, float b, float& c_lo, float& c_hi) {
    const f2_t d = f2_fma(x2, f2_make(-1.0f, -1.0f), f2_make(b, b));
    const f2_t k = f2_add(f2_add(d, f2_make(NRC_RINT_MAGIC, NRC_RINT_MAGIC)), f2_make(-NRC_RINT_MAGIC, -NRC_RINT_MAGIC));
    const f2_t r = f2_fma(k, f2_make(-1.0f, -1.0f), d);
    const f2_t q2 = f2_mul(r, r);
    const f2_t q4 = f2_mul(q2, q2);
    const float k1 = 16.0f * -(2.0f / 3.0f), k2 = 256.0f * 0.2f;
    const f2_t inner = f2_fma(q4, f2_make(k2, k2), f2_fma(q2, f2_make(k1, k1), f2_make(1.0f, 1.0f)));
    const f2_t a = f2_mul(r, f2_make(3.75f, 3.75f));
    const f2_t kp = f2_add(k, f2_make(1.0f, 1.0f));
    float a0, a1, i0, i1, kp0, kp1;
    f2_split(a, a0, a1); f2_split(inner, i0, i1); f2_split(kp, kp0, kp1);
    c_lo = fminf(fmaxf(kp0 + __saturatef(fmaf(a0, i0, 0.5f)), 0.0f), 3.0f);
    c_hi = fminf(fmaxf(kp1 + __saturatef(fmaf(a1, i1, 0.5f)), 0.0f), 3.0f);
}
// 4 bins of two dimensions (lanes of x2) -> ya / yb.
template <bool WRAP_LAST>
__device__ __forceinline__ void oneblob4_x2(f2_t x2, float (&ya)[4], float (&yb)[4]) {
    float ca[5], cb[5];
    blob_cdf3_x2(x2, 0.0f, ca[0], cb[0]);
    blob_cdf3_x2(x2, 0.25f, ca[1], cb[1]);
    blob_cdf3_x2(x2, 0.5f, ca[2], cb[2]);
    blob_cdf3_x2(x2, 0.75f, ca[3], cb[3]);
    if (WRAP_LAST) { ca[4] = ca[0] + 1.0f; cb[4] = cb[0] + 1.0f; }
    else blob_cdf3_x2(x2, 1.0f, ca[4], cb[4]);
#pragma unroll
    for (int i = 0; i < 4; ++i) { ya[i] = ca[i + 1] - ca[i]; yb[i] = cb[i + 1] - cb[i]; }
}
__device__ __forceinline__ float triwave(float x, int f) {
    // scalbnf(x, f-1) is exact, so one fma reproduces  x*2^(f-1) + 0.25 f  with a single rounding
    const float v = fmaf(x, __int_as_float((126 + f) << 23), 0.25f * (float)f);
    const float r = v - floorf(v);
    return fmaf(fabsf(r - 0.5f), 4.0f, -1.0f);
}

// Folded tri-wave: with d = v - rint(v) in [-1/2, 1/2] (exact),  |frac(v) - 1/2| = 1/2 - |d|,  so  y = 1 - 4|d|  with ONE
// rounding (the reference rounds frac(v) - 1/2 first: a 2^-26 difference that the fp16 rounding of the feature hides in all
// but ~1e-4 of the cases, where the feature moves by one fp16 ulp).  rint through the magic add: |v| < 2^22, i.e. scene
// coordinates below 4096 / 0.005 units.  No XU-pipe instruction (floorf is FRND, 8 issue cycles per warp).
__device__ __forceinline__ float triwave_fold(float x, int f) {
    const float v = fmaf(x, __int_as_float((126 + f) << 23), 0.25f * (float)f);
    const float n = (v + NRC_RINT_MAGIC) - NRC_RINT_MAGIC;
    return fmaf(fabsf(v - n), -4.0f, 1.0f);
}
// Tri-wave of x and y -> 12 packed half2 words (word d*6 + f/2 holds frequencies f, f + 1 of coordinate d).
__device__ __forceinline__ void triwave_xy(float x, float y, uint32_t* e) {
    float t[24];
#if NRC_ENC_VARIANT == 0
#pragma unroll
    for (int f = 0; f < 12; ++f) { t[f] = triwave(x, f); t[12 + f] = triwave(y, f); }
#elif NRC_ENC_VARIANT == 1
#pragma unroll
    for (int f = 0; f < 12; ++f) { t[f] = triwave_fold(x, f); t[12 + f] = triwave_fold(y, f); }
#else
    // lanes = (x, y): every constant is a broadcast
    const f2_t xy = f2_make(x, y);
    const f2_t mp = f2_make(NRC_RINT_MAGIC, NRC_RINT_MAGIC), mn = f2_make(-NRC_RINT_MAGIC, -NRC_RINT_MAGIC), neg1 = f2_make(-1.0f, -1.0f);
#pragma unroll
    for (int f = 0; f < 12; ++f) {
        const float s = __int_as_float((126 + f) << 23), o = 0.25f * (float)f;
        const f2_t v = f2_fma(xy, f2_make(s, s), f2_make(o, o));
        const f2_t d = f2_fma(f2_add(f2_add(v, mp), mn), neg1, v);
        float d0, d1;
        f2_split(d, d0, d1);
        t[f] = fmaf(fabsf(d0), -4.0f, 1.0f);
        t[12 + f] = fmaf(fabsf(d1), -4.0f, 1.0f);
    }
#endif
#pragma unroll
    for (int i = 0; i < 12; ++i) e[i] = pack_h2(t[2 * i], t[2 * i + 1]);
}
// Tri-wave of z -> 6 packed half2 words.
__device__ __forceinline__ void triwave_z(float z, uint32_t* e) {
    float t[12];
#if NRC_ENC_VARIANT == 0
#pragma unroll
    for (int f = 0; f < 12; ++f) t[f] = triwave(z, f);
#elif NRC_ENC_VARIANT == 1
#pragma unroll
    for (int f = 0; f < 12; ++f) t[f] = triwave_fold(z, f);
#else
    // lanes = frequencies (f, f + 1)
    const f2_t zz = f2_make(z, z);
    const f2_t mp = f2_make(NRC_RINT_MAGIC, NRC_RINT_MAGIC), mn = f2_make(-NRC_RINT_MAGIC, -NRC_RINT_MAGIC), neg1 = f2_make(-1.0f, -1.0f);
#pragma unroll
    for (int f = 0; f < 12; f += 2) {
        const f2_t v = f2_fma(zz, f2_make(__int_as_float((126 + f) << 23), __int_as_float((127 + f) << 23)),
                              f2_make(0.25f * (float)f, 0.25f * (float)(f + 1)));
        const f2_t d = f2_fma(f2_add(f2_add(v, mp), mn), neg1, v);
        float d0, d1;
        f2_split(d, d0, d1);
        t[f] = fmaf(fabsf(d0), -4.0f, 1.0f);
        t[f + 1] = fmaf(fabsf(d1), -4.0f, 1.0f);
    }
#endif
#pragma unroll
    for (int i = 0; i < 6; ++i) e[i] = pack_h2(t[2 * i], t[2 * i + 1]);
}
// Tri-wave of the three position coordinates -> 18 packed half2 words (j = d*12 + f).
__device__ __forceinline__ void triwave_xyz(const float (&q)[15], uint32_t* e) {
    triwave_xy(q[0], q[1], e);
    triwave_z(q[2], e + 12);
}
// One-blob of two neighbouring dimensions -> 4 packed half2 words.
template <bool WRAP_LAST>
__device__ __forceinline__ void oneblob_pair(float xa, float xb, uint32_t* e) {
    float ya[4], yb[4];
#if NRC_ENC_VARIANT >= 3
    oneblob4_x2<WRAP_LAST>(f2_make(xa, xb), ya, yb);
#else
    oneblob4<WRAP_LAST>(xa, ya);
    oneblob4<WRAP_LAST>(xb, yb);
#endif
    e[0] = pack_h2(ya[0], ya[1]); e[1] = pack_h2(ya[2], ya[3]);
    e[2] = pack_h2(yb[0], yb[1]); e[3] = pack_h2(yb[2], yb[3]);
}

// The Frequency encoding in five independent slices that write their half2 words straight into the K-major A image of
// the input layer (a_row = this query's row; 16-byte chunk c = words 4c .. 4c+3 at + c*2048): the query kernel runs
// slice k of the NEXT tile while the tensor pipe works on layer k + 1 of the current one.  Words: tri-wave x,y 0..11 |
// tri-wave z 12..17 | one-blob (3,4) 18..21 | (5,6) 22..25 | (7,8) 26..29 | albedos 30..32 | 1.0 pad 33..39.
__device__ __forceinline__ void a_store2(uint8_t* a_row, int w, uint32_t v0, uint32_t v1) {   // words w, w + 1 (w even)
    *reinterpret_cast<uint2*>(a_row + (w >> 2) * 2048 + (w & 3) * 4) = make_uint2(v0, v1);
}
__device__ __forceinline__ void a_store4(uint8_t* a_row, int w, const uint32_t* v) {          // words w .. w + 3 (w % 4 == 0)
    *reinterpret_cast<uint4*>(a_row + (w >> 2) * 2048) = make_uint4(v[0], v[1], v[2], v[3]);
}
template <int S>
__device__ __forceinline__ void encode_frequency_slice(const float (&q)[15], uint8_t* a_row) {
    if constexpr (S == 0) {
        uint32_t e[12];
        triwave_xy(q[0], q[1], e);
        a_store4(a_row, 0, e); a_store4(a_row, 4, e + 4); a_store4(a_row, 8, e + 8);
    } else if constexpr (S == 1) {
        uint32_t e[6];
        triwave_z(q[2], e);
        a_store4(a_row, 12, e);
        a_store2(a_row, 16, e[4], e[5]);
        a_store2(a_row, 30, pack_h2(q[9], q[10]), pack_h2(q[11], q[12]));
        const uint32_t one = 0x3C003C00u;   // (1.0h, 1.0h)
        const uint32_t w8[4] = {pack_h2(q[13], q[14]), one, one, one}, w9[4] = {one, one, one, one};
        a_store4(a_row, 32, w8);
        a_store4(a_row, 36, w9);
    } else {
        uint32_t e[4];
        oneblob_pair<true>(q[3 + 2 * (S - 2)], q[4 + 2 * (S - 2)], e);
        a_store2(a_row, 18 + 4 * (S - 2), e[0], e[1]);
        a_store2(a_row, 20 + 4 * (S - 2), e[2], e[3]);
    }
}

// Frequency config: 15 raw floats -> 80 fp16 features packed as 40 half2 words (lo = even feature).
__device__ __forceinline__ void encode_frequency(const float (&q)[15], uint32_t (&e)[40]) {
    triwave_xyz(q, e);
#pragma unroll
    for (int d = 0; d < 6; d += 2) oneblob_pair<true>(q[3 + d], q[4 + d], e + 18 + 2 * d);
    e[30] = pack_h2(q[9], q[10]);
    e[31] = pack_h2(q[11], q[12]);
    e[32] = pack_h2(q[13], q[14]);
#pragma unroll
    for (int i = 33; i < 40; ++i) e[i] = 0x3C003C00u;  // (1.0h, 1.0h)
}

// One level of the hash grid: trilinear interpolation of half2 features, accumulated with half2 fma
// exactly like the reference (grid.h:146-163: result = fma((T)weight, grid_val, result); corner c = dx + 2 dy + 4 dz,
// weight = ((1 * fx) * fy) * fz).
// NC = true: read-only (ld.global.nc) loads -- query kernel; false: coherent loads -- the training kernel updates the
// table later in the same launch.
// The per-axis index terms are computed once per level (2 values per axis, already masked: & distributes over ^), so a
// corner costs one 3-input LOP3 (hashed) or one IADD3 + mask (dense) instead of two multiplies and two xors.
// Which levels are hashed follows the reference's grid_index literally (GridMeta, filled by the host): besides the truly
// dense coarse levels, fine levels whose stride res^2 wraps to 0 in uint32 (res >= 2^16 with a table >= res entries) come
// out of that loop un-hashed as x + y*res.
template <bool NC = true, int LVL = 0>
__device__ __forceinline__ uint32_t grid_level(const __half2* table, uint32_t size_mask, uint32_t sy, uint32_t sz,
                                               float scale, bool dense, float x, float y, float z) {
    float p[3] = {fmaf(scale, x, 0.5f), fmaf(scale, y, 0.5f), fmaf(scale, z, 0.5f)};
    uint32_t g[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float fl = floorf(p[d]);
        g[d] = (uint32_t)(int)fl;
        p[d] -= fl;
    }
    const float fx[2] = {1.0f - p[0], p[0]}, fy[2] = {1.0f - p[1], p[1]}, fz[2] = {1.0f - p[2], p[2]};
    const float wxy[4] = {fx[0] * fy[0], fx[1] * fy[0], fx[0] * fy[1], fx[1] * fy[1]};
    uint32_t idx[8];
    if (dense) {
        const uint32_t cy[2] = {g[1] * sy, g[1] * sy + sy}, cz[2] = {g[2] * sz, g[2] * sz + sz};
#pragma unroll
        for (int c = 0; c < 8; ++c) idx[c] = (g[0] + (c & 1) + cy[(c >> 1) & 1] + cz[c >> 2]) & size_mask;
    } else {
        const uint32_t cx[2] = {g[0] & size_mask, (g[0] + 1) & size_mask};
        const uint32_t hy = g[1] * 2654435761u, hz = g[2] * 805459861u;
        const uint32_t cy[2] = {hy & size_mask, (hy + 2654435761u) & size_mask};
        const uint32_t cz[2] = {hz & size_mask, (hz + 805459861u) & size_mask};
#pragma unroll
        for (int c = 0; c < 8; ++c) idx[c] = cx[c & 1] ^ cy[(c >> 1) & 1] ^ cz[c >> 2];
    }
    __half2 acc = __float2half2_rn(0.0f);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float w = wxy[c & 3] * fz[c >> 2];
        __half2 v;
        if constexpr (NC) v = __ldg(table + idx[c]);
        else v = table[idx[c]];
        acc = __hfma2(__float2half2_rn(w), v, acc);
    }
    return *reinterpret_cast<uint32_t*>(&acc);
}

// ---- Pipelined gather for the query kernel ------------------------------------------------------------------------------
// Same arithmetic as grid_level above (weights in fp32, fp16 FMA chain in corner order c = dx + 2 dy + 4 dz), restructured:
//  * issue / consume are split so that the loads of level l+1 are in flight while the FMA chain of level l runs
//    (GridFetch holds what is pending);
//  * floor() goes through the round-DOWN magic add: t = p + 1.5*2^23 rounded toward -inf holds floor(p) in its low mantissa
//    bits (exact for |p| < 2^22, i.e. |x| < 8 at the finest level) -- no XU-pipe instruction (FRND / F2I) is left.  The
//    constant 0x4B400000 the bit pattern carries is never subtracted per axis: its low 22 bits are zero, so it vanishes
//    under the table mask (tables <= 2^22 entries) and folds into the addend of the per-axis multiply otherwise;
//  * one IMAD.WIDE per address (the table base is kept in a vector register pair on purpose: with a uniform base ptxas
//    emits LEA + LEA.HI.X, two issue slots per gather), two weights per F2FP (the FMA selects its half with .H0_H0 / .H1_H1).
// Measured alternatives, kept as A/B builds (profiles/r02_ab_variants.json, ncu: profiles/r02_infer_hash_*): NRC_HASH_PAIR=1
// fetches x-neighbours as one aligned 8-byte pair where the first corner's index is even (-25 % L1 sectors, but +25 %
// instructions for the selects: slower on the pixel-ordered frame, faster on incoherent batches); NRC_HASH_SMEM0=1 stages
// level 0 (16 KB) in shared memory (costs L1 capacity the fine levels need: slower on both).
#ifndef NRC_HASH_PAIR
#define NRC_HASH_PAIR 0
#endif
struct GridFetch {
    float f[3];                 // position inside the cell
#if NRC_HASH_PAIR
    uint32_t pa[4], pb[4];      // low / high word of the aligned pair that holds corner (0, dy, dz)
    uint32_t c1[4];             // corner (1, dy, dz) where it is not the pair's other word (lanes with sep set)
    uint32_t lo;                // bit j: index of corner (0, dy, dz) is odd (it is the pair's HIGH word)
    uint32_t sep;               // bit j: corner (1, dy, dz) is NOT the other word of the pair
#else
    uint32_t v[8];
#endif
};
__device__ __forceinline__ uint64_t entry_addr(uint64_t table, uint32_t idx) {     // table + 4 * idx: one IMAD.WIDE.U32
    uint64_t a;
    asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(a) : "r"(idx), "l"(table));
    return a;
}
// SMEM: the level's table has been staged in shared memory (smem_table = its 32-bit shared address).
// NC: read-only (ld.global.nc) loads -- query kernel; false: plain loads -- the training kernel updates the table later in the
// same launch.
template <bool SMEM, bool NC = true>
__device__ __forceinline__ void grid_issue(GridFetch& F, const __half2* table_ptr, uint32_t smem_table, uint32_t size_mask, uint32_t sy,
                                           uint32_t sz, float scale, bool dense, float x, float y, float z) {
    constexpr uint32_t kMagicBits = 0x4B400000u;                       // bit pattern of 1.5 * 2^23
    const float p[3] = {fmaf(scale, x, 0.5f), fmaf(scale, y, 0.5f), fmaf(scale, z, 0.5f)};
    uint32_t tb[3];                                                    // kMagicBits + floor(p) (two's complement in the low bits)
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float t = __fadd_rd(p[d], NRC_RINT_MAGIC);               // floor(p) + 1.5 * 2^23, exact
        tb[d] = __float_as_uint(t);
        F.f[d] = p[d] - (t - NRC_RINT_MAGIC);
    }
    uint64_t table = reinterpret_cast<uint64_t>(table_ptr);
    asm volatile("" : "+l"(table));                                    // keep the base in vector registers (see above)
    uint32_t i0[4], i1[4];
    if (dense) {
        // (g0 + g1*sy + g2*sz) with g = tb - kMagicBits: the three constants collapse into one addend
        const uint32_t base = tb[0] + tb[1] * sy + tb[2] * sz - kMagicBits * (1u + sy + sz);
        const uint32_t cyz[4] = {base, base + sy, base + sz, base + sy + sz};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            i0[j] = cyz[j] & size_mask;
            i1[j] = (cyz[j] + 1) & size_mask;
        }
    } else {
        // masked per-axis terms (& distributes over ^).  x: prime 1; the magic constant's low 22 bits are zero, so it drops
        // out under masks below 2^22 and is subtracted only for larger tables.  y, z: folded into the multiply's addend.
        const uint32_t g0 = size_mask < (1u << 22) ? tb[0] : tb[0] - kMagicBits;
        const uint32_t cx[2] = {g0 & size_mask, (g0 + 1) & size_mask};
        const uint32_t hy = tb[1] * 2654435761u - kMagicBits * 2654435761u, hz = tb[2] * 805459861u - kMagicBits * 805459861u;
        const uint32_t cy[2] = {hy & size_mask, (hy + 2654435761u) & size_mask};
        const uint32_t cz[2] = {hz & size_mask, (hz + 805459861u) & size_mask};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            i0[j] = cx[0] ^ cy[j & 1] ^ cz[j >> 1];
            i1[j] = cx[1] ^ cy[j & 1] ^ cz[j >> 1];
        }
    }
#if NRC_HASH_PAIR
    // bit 0 of the four first-corner indices: lo = sum (i0[j] & 1) << j
    F.lo = (i0[0] & 1u) | ((i0[1] & 1u) << 1) | ((i0[2] & 1u) << 2) | ((i0[3] & 1u) << 3);
    // hashed: the x neighbour is the other word of the pair exactly when x is even (then idx(x+1) = idx(x) ^ 1): one bit for the
    // four pairs; dense: when the first index is even.
    F.sep = dense ? F.lo : ((tb[0] & 1u) ? 0xFu : 0u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t need = (F.sep >> j) & 1u;
        F.c1[j] = 0u;
        if constexpr (SMEM) {
            const uint32_t a0 = smem_table + (i0[j] & ~1u) * 4u, a1 = smem_table + i1[j] * 4u;
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(F.pa[j]), "=r"(F.pb[j]) : "r"(a0));
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p ld.shared.b32 %0, [%1];\n\t}" : "+r"(F.c1[j]) : "r"(a1), "r"(need));
        } else {
            asm volatile("ld.global.nc.v2.b32 {%0, %1}, [%2];" : "=r"(F.pa[j]), "=r"(F.pb[j]) : "l"(entry_addr(table, i0[j] & ~1u)));
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p ld.global.nc.b32 %0, [%1];\n\t}" : "+r"(F.c1[j]) : "l"(entry_addr(table, i1[j])), "r"(need));
        }
    }
#else
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if constexpr (SMEM) {
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(F.v[2 * j]) : "r"(smem_table + i0[j] * 4u));
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(F.v[2 * j + 1]) : "r"(smem_table + i1[j] * 4u));
        } else {
            if constexpr (NC) {
                asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(F.v[2 * j]) : "l"(entry_addr(table, i0[j])));
                asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(F.v[2 * j + 1]) : "l"(entry_addr(table, i1[j])));
            } else {
                asm volatile("ld.global.b32 %0, [%1];" : "=r"(F.v[2 * j]) : "l"(entry_addr(table, i0[j])) : "memory");
                asm volatile("ld.global.b32 %0, [%1];" : "=r"(F.v[2 * j + 1]) : "l"(entry_addr(table, i1[j])) : "memory");
            }
        }
    }
#endif
}
// The level's two features: result = fma((half)w_c, v_c, result) over the corners in the reference's order (grid.h:146-163),
// w_c = ((1 * wx) * wy) * wz in fp32.  Two weights are converted per F2FP; the FMA reads its half through the .H0_H0 /
// .H1_H1 operand selectors.
__device__ __forceinline__ uint32_t grid_consume(const GridFetch& F) {
    const float fx[2] = {1.0f - F.f[0], F.f[0]}, fy[2] = {1.0f - F.f[1], F.f[1]}, fz[2] = {1.0f - F.f[2], F.f[2]};
    const float wxy[4] = {fx[0] * fy[0], fx[1] * fy[0], fx[0] * fy[1], fx[1] * fy[1]};
    __half2 acc = __float2half2_rn(0.0f);
#pragma unroll
    for (int j = 0; j < 4; ++j) {          // j = dy + 2 dz: corners c = 2j (dx = 0) and 2j + 1 (dx = 1)
#if NRC_HASH_PAIR
        const bool hi = (F.lo >> j) & 1u, sep = (F.sep >> j) & 1u;
        const uint32_t v0 = hi ? F.pb[j] : F.pa[j];
        const uint32_t other = hi ? F.pa[j] : F.pb[j];
        const uint32_t v1 = sep ? F.c1[j] : other;
#else
        const uint32_t v0 = F.v[2 * j], v1 = F.v[2 * j + 1];
#endif
        const uint32_t wp = pack_h2(wxy[2 * (j & 1)] * fz[j >> 1], wxy[2 * (j & 1) + 1] * fz[j >> 1]);
        const __half2 w2 = *reinterpret_cast<const __half2*>(&wp);
        acc = __hfma2(__low2half2(w2), *reinterpret_cast<const __half2*>(&v0), acc);
        acc = __hfma2(__high2half2(w2), *reinterpret_cast<const __half2*>(&v1), acc);
    }
    return *reinterpret_cast<uint32_t*>(&acc);
}
template <int L, bool SMEM0, bool NC = true>
__device__ __forceinline__ void grid_issue_level(GridFetch& F, const float (&q)[15], const __half2* grid, uint32_t smem_level0, const GridMeta& gm) {
    const uint32_t size = gm.offset[L + 1] - gm.offset[L];
    grid_issue<(L == 0 && SMEM0), NC>(F, grid + gm.offset[L], smem_level0, size - 1, gm.stride_y[L], gm.stride_z[L], gm.scale[L],
                                  (gm.dense_mask >> L) & 1, q[0], q[1], q[2]);
}

// Levels [L0, L1) -> e[L0 .. L1), the level index a compile-time constant.
template <bool NC, int L0, int L1, int N>
__device__ __forceinline__ void grid_levels(const float (&q)[15], const __half2* grid, const GridMeta& gm, uint32_t (&e)[N]) {
    if constexpr (L0 < L1) {
        const uint32_t size = gm.offset[L0 + 1] - gm.offset[L0];
        e[L0] = grid_level<NC, L0>(grid + gm.offset[L0], size - 1, gm.stride_y[L0], gm.stride_z[L0], gm.scale[L0], (gm.dense_mask >> L0) & 1,
                                   q[0], q[1], q[2]);
        grid_levels<NC, L0 + 1, L1>(q, grid, gm, e);
    }
}

// Hash config: 15 raw floats -> 64 fp16 features packed as 32 half2 words. Level tables are powers of two
// (NRCNetworkConfigs.h:98-104: 16^3 = 4096 entries at level 0, 2^log2_hashmap_size above), so `% size` is a mask.
__device__ __forceinline__ void encode_hash(const float (&q)[15], const __half2* __restrict__ grid,
                                            const GridMeta& gm, uint32_t (&e)[32]) {
    grid_levels<true, 0, 16>(q, grid, gm, e);
#pragma unroll
    for (int d = 0; d < 6; d += 2) oneblob_pair<false>(q[3 + d], q[4 + d], e + 16 + 2 * d);
    e[28] = pack_h2(q[9], q[10]);
    e[29] = pack_h2(q[11], q[12]);
    e[30] = pack_h2(q[13], q[14]);
    e[31] = 0x3C003C00u;
}

// ---- Half encoders for the training kernel: two threads share one sample, each produces half of the 16-byte
// feature chunks (8 features per chunk).  HF = 0: chunks [0, E/16); HF = 1: chunks [E/16, E/8).
// Frequency (10 chunks): HF0 = tri-wave 36 + one-blob of dim 3 (features 0..39); HF1 = one-blob dims 4..8, albedos, pad.
template <int HF>
__device__ __forceinline__ void encode_frequency_half(const float (&q)[15], uint32_t (&e)[20]) {
    if constexpr (HF == 0) {
        triwave_xyz(q, e);   // same instruction sequence as the query kernel: training and queries see identical features
        float y[4];
        oneblob4<true>(q[3], y);
        e[18] = pack_h2(y[0], y[1]);
        e[19] = pack_h2(y[2], y[3]);
    } else {
#pragma unroll
        for (int d = 1; d < 6; ++d) {
            float y[4];
            oneblob4<true>(q[3 + d], y);
            e[2 * d - 2] = pack_h2(y[0], y[1]);
            e[2 * d - 1] = pack_h2(y[2], y[3]);
        }
        e[10] = pack_h2(q[9], q[10]);
        e[11] = pack_h2(q[11], q[12]);
        e[12] = pack_h2(q[13], q[14]);
#pragma unroll
        for (int i = 13; i < 20; ++i) e[i] = 0x3C003C00u;
    }
}
// Hash, training kernel: the two threads of a sample split the gather -- thread HF runs levels 8 HF .. 8 HF + 7 through the
// pipelined gather (plain loads) and a part of the cheap features.  Output: four 16-byte chunks of the K-major feature row,
// chunk ids in `ids` (chunk c = features 8c .. 8c+7: grid levels 4c .. 4c+3 for c < 4, one-blob pairs for c = 4..6, albedos + pad 7).
//   HF0: chunks 0, 1 (levels 0..7), 4, 5 (one-blob of dims 0..3)      HF1: chunks 2, 3 (levels 8..15), 6 (dims 4, 5), 7
#ifndef NRCB_HASH_TRAIN_DEPTH
#define NRCB_HASH_TRAIN_DEPTH 2     // levels of gathers in flight per thread in the training encode; 4 and 8 measured: no change (6.1 / 6.3 / 6.6 us:
                                    // the encode is bound by L2 -> SM sector traffic with a cold L1, not by the load latency chain) and more registers
#endif
template <int L0, int L, int D>
__device__ __forceinline__ void hash_half_levels(GridFetch (&F)[D], uint32_t (&e)[16], const float (&q)[15], const __half2* grid, const GridMeta& gm) {
    if constexpr (L < L0 + 8) {
        if constexpr (L + D - 1 < L0 + 8) grid_issue_level<L + D - 1, false, false>(F[(L + D - 1) % D], q, grid, 0u, gm);
        e[L - L0] = grid_consume(F[L % D]);
        hash_half_levels<L0, L + 1, D>(F, e, q, grid, gm);
    }
}
template <int L0, int L, int D>
__device__ __forceinline__ void hash_half_prologue(GridFetch (&F)[D], const float (&q)[15], const __half2* grid, const GridMeta& gm) {
    if constexpr (L < L0 + D - 1) {
        grid_issue_level<L, false, false>(F[L % D], q, grid, 0u, gm);
        hash_half_prologue<L0, L + 1, D>(F, q, grid, gm);
    }
}
template <int HF>
__device__ __forceinline__ void encode_hash_split(const float (&q)[15], const __half2* grid, const GridMeta& gm, uint32_t (&e)[16], int (&ids)[4]) {
    constexpr int D = NRCB_HASH_TRAIN_DEPTH;
    static_assert(D >= 2 && D <= 8, "levels in flight");
    GridFetch F[D];
    hash_half_prologue<8 * HF, 8 * HF, D>(F, q, grid, gm);
    hash_half_levels<8 * HF, 8 * HF, D>(F, e, q, grid, gm);
    if constexpr (HF == 0) {
        ids[0] = 0; ids[1] = 1; ids[2] = 4; ids[3] = 5;
        oneblob_pair<false>(q[3], q[4], e + 8);
        oneblob_pair<false>(q[5], q[6], e + 12);
    } else {
        ids[0] = 2; ids[1] = 3; ids[2] = 6; ids[3] = 7;
        oneblob_pair<false>(q[7], q[8], e + 8);
        e[12] = pack_h2(q[9], q[10]);
        e[13] = pack_h2(q[11], q[12]);
        e[14] = pack_h2(q[13], q[14]);
        e[15] = 0x3C003C00u;
    }
}
// Round-1 form (kept for A/B builds): HF0 = the 16 grid levels (features 0..31); HF1 = one-blob, albedos, pad (features 32..63).
template <int HF>
__device__ __forceinline__ void encode_hash_half(const float (&q)[15], const __half2* grid,
                                                 const GridMeta& gm, uint32_t (&e)[16]) {
    if constexpr (HF == 0) {
        grid_levels<false, 0, 16>(q, grid, gm, e);
    } else {
#pragma unroll
        for (int d = 0; d < 6; ++d) {
            float y[4];
            oneblob4<false>(q[3 + d], y);
            e[2 * d] = pack_h2(y[0], y[1]);
            e[2 * d + 1] = pack_h2(y[2], y[3]);
        }
        e[12] = pack_h2(q[9], q[10]);
        e[13] = pack_h2(q[11], q[12]);
        e[14] = pack_h2(q[13], q[14]);
        e[15] = 0x3C003C00u;
    }
}

}  // namespace nrcb
