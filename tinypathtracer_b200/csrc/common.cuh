// common.cuh — device math, RNG and error plumbing shared by the kernels of libtpt_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <string>

#include "tpt.h"

namespace tpt {

// ------------------------------------------------------------------------------------------------
// errors (the C ABI returns codes; message kept per host thread)
// ------------------------------------------------------------------------------------------------
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define TPT_CUDA(call)                                                          \
    do {                                                                        \
        cudaError_t e__ = (call);                                               \
        if (e__ != cudaSuccess) return ::tpt::cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)
#define TPT_CHECK(expr)                 \
    do {                                \
        int rc__ = (expr);              \
        if (rc__ != TPT_OK) return rc__; \
    } while (0)
#define TPT_REQUIRE(cond, msg)              \
    do {                                    \
        if (!(cond)) {                      \
            ::tpt::set_error(msg);          \
            return TPT_ERR_INVALID;         \
        }                                   \
    } while (0)

constexpr int kNumSMs = 148;        // B200: 2 dies x 74 SMs
constexpr float kPi = 3.141592653589793f;
constexpr float kDelta = 2e-4f;     // reference MathConst::Delta (include/math/vec.h:70)
constexpr int kAccumFracBits = 36;  // fixed-point accumulation: value * 2^36 in int64

// ------------------------------------------------------------------------------------------------
// float3 helpers
// ------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__host__ __device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__host__ __device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ float3 cross(float3 a, float3 b) {
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__host__ __device__ __forceinline__ float3 fmin3(float3 a, float3 b) { return f3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
__host__ __device__ __forceinline__ float3 fmax3(float3 a, float3 b) { return f3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

// The reference normalises every direction and shading normal with the Quake reciprocal square
// root and ONE Newton step (include/math/vec.h:25-38,155): lengths are off by up to 0.175 %, and hit
// distances are measured in units of that non-unit direction.  t-parity at 1e-5 requires the same
// function, so it is part of the contract, not an approximation of ours.
__host__ __device__ __forceinline__ float ref_frsqrt(float num) {
    float x2 = num * 0.5f;
#ifdef __CUDA_ARCH__
    int i = __float_as_int(num);
    i = 0x5f3759df - (i >> 1);
    float y = __int_as_float(i);
#else
    int i;
    float y = num;
    memcpy(&i, &y, 4);
    i = 0x5f3759df - (i >> 1);
    memcpy(&y, &i, 4);
#endif
    return y * (1.5f - (x2 * y * y));
}
__host__ __device__ __forceinline__ float3 ref_normalize(float3 v) { return v * ref_frsqrt(dot(v, v)); }

// ------------------------------------------------------------------------------------------------
// Exactly-rounded arithmetic for the geometric core.  The quantities the parity bars are stated on
// (world vertices, primary rays, hit distance t, barycentrics, primitive id) are computed with one
// IEEE rounding per operation in the reference's operand order — no FMA contraction — so that they
// are BIT-IDENTICAL to the reference source evaluated op by op (which is what the CPU oracle and a
// host compile of the reference's CUDA_CALLABLE code produce).  __fmul_rn/__fadd_rn are never fused
// by nvcc.  Shading arithmetic (BSDF, lights, throughput) is left to the compiler: it is only
// compared statistically.
// ------------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ float xmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float3 xsub3(float3 a, float3 b) { return f3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
__device__ __forceinline__ float xdot(float3 a, float3 b) {   // vec.h:151, left-associated
    return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z));
}
__device__ __forceinline__ float3 xcross(float3 a, float3 b) {   // vec.h:152
    return f3(xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)), xsub(xmul(a.x, b.y), xmul(a.y, b.x)));
}
__device__ __forceinline__ float xfrsqrt(float num) {   // vec.h:25-38
    const float x2 = xmul(num, 0.5f);
    float y = __int_as_float(0x5f3759df - (__float_as_int(num) >> 1));
    return xmul(y, xsub(1.5f, xmul(xmul(x2, y), y)));
}
__device__ __forceinline__ float3 xnormalize(float3 v) {   // vec.h:155: v * frsqrt(norm2)
    const float s = xfrsqrt(xdot(v, v));
    return f3(xmul(v.x, s), xmul(v.y, s), xmul(v.z, s));
}
// MatrixVectorMultiply (mat.h:37-51): res[i] = (((0 + m[0][i]*v0) + m[1][i]*v1) + m[2][i]*v2) + m[3][i]*v3
__device__ __forceinline__ float3 xmat4_point(const float* m, float3 p) {
    return f3(xadd(xadd(xadd(xmul(m[0], p.x), xmul(m[4], p.y)), xmul(m[8], p.z)), m[12]),
              xadd(xadd(xadd(xmul(m[1], p.x), xmul(m[5], p.y)), xmul(m[9], p.z)), m[13]),
              xadd(xadd(xadd(xmul(m[2], p.x), xmul(m[6], p.y)), xmul(m[10], p.z)), m[14]));
}
__device__ __forceinline__ float3 xmat4_dir(const float* m, float3 d) {   // w = 0: the last term adds +-0
    return f3(xadd(xadd(xadd(xmul(m[0], d.x), xmul(m[4], d.y)), xmul(m[8], d.z)), xmul(m[12], 0.0f)),
              xadd(xadd(xadd(xmul(m[1], d.x), xmul(m[5], d.y)), xmul(m[9], d.z)), xmul(m[13], 0.0f)),
              xadd(xadd(xadd(xmul(m[2], d.x), xmul(m[6], d.y)), xmul(m[10], d.z)), xmul(m[14], 0.0f)));
}
#endif

// column-major Mat4 (reference include/math/mat.h:37-51)
__host__ __device__ __forceinline__ float3 mat4_point(const float* m, float3 p) {
    return f3(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
              m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]);
}
__host__ __device__ __forceinline__ float3 mat4_dir(const float* m, float3 d) {
    return f3(m[0] * d.x + m[4] * d.y + m[8] * d.z, m[1] * d.x + m[5] * d.y + m[9] * d.z, m[2] * d.x + m[6] * d.y + m[10] * d.z);
}

// ------------------------------------------------------------------------------------------------
// RNG — counter-based, stateless.  Spec (DESIGN.md "RNG spec"; oracle/oracle.cpp restates it):
//   key(seed, pixel, sample) = mix64( mix64(seed + G) ^ (pixel << 32 | sample) ),  G = 0x9E3779B97F4A7C15
//   pair(key, j)             = z_j -> a = u01(hi32 z_j), b = u01(lo32 z_j);  z_0 = key,  z_j = mix64( key + j * G ) for j >= 1
//   u01(x)                   = ((x >> 8) + 1) * 2^-24   in (0, 1]   (the range of curand_uniform)
//   mix64 = the SplitMix64 finaliser.
// Pair 0 is the pixel jitter; at path vertex d pair 1+2d drives the BSDF sample (dielectric: a is the
// Fresnel coin; diffuse: a -> phi, b -> cos^2) and pair 2+2d the extra "direct" cosine ray.
// Replaces the per-pixel cuRAND XORWOW state in global memory (include/sampler.h:12-15,
// src/path_tracer.cu:34-40): no state buffer, and a sample's value does not depend on which GPU,
// kernel or thread evaluates it.
// ------------------------------------------------------------------------------------------------
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t rng_seed_key(uint64_t seed) { return mix64(seed + kGolden); }
__host__ __device__ __forceinline__ uint64_t rng_path_key(uint64_t seedKey, uint32_t pixel, uint32_t sample) {
    return mix64(seedKey ^ (((uint64_t)pixel << 32) | (uint64_t)sample));
}
__host__ __device__ __forceinline__ float rng_u01(uint32_t x) { return (float)((x >> 8) + 1u) * (1.0f / 16777216.0f); }

// Spectrum::toUChar, include/material.h:74-81: (unsigned char) clamp(255 * L / spp, 0, 255) by truncation; one fixed-point
// accumulator word (2^-36 units) -> mean radiance.  Shared by k_resolve (render.cu) and k_xchg_resolve (comm.cu): same bytes.
__device__ __forceinline__ unsigned char to_uchar(float v) {
    const float x = v * 255.0f;
    const float c = fmaxf((x < 255.0f) ? x : 255.0f, 0.0f);   // min(x,255) with the reference's NaN -> 255
    return (unsigned char)c;
}
__device__ __forceinline__ float accum_to_radiance(unsigned long long a, float invSpp) {
    return (float)((double)(long long)a * (1.0 / 68719476736.0)) * invSpp;
}
// the 64 raw bits behind pair j (env.cuh picks an alias-table bucket from all 32 high bits: a 24-bit float cannot address a
// 4096 x 2048 map's 2^23 texels AND leave bits for anything else)
__host__ __device__ __forceinline__ uint64_t rng_bits(uint64_t key, uint32_t j) {
    // pair 0 (the pixel jitter, drawn for every camera sample) uses the already-mixed key itself: one mix64 less per sample
    return j == 0u ? key : mix64(key + (uint64_t)j * kGolden);
}
__host__ __device__ __forceinline__ void rng_pair(uint64_t key, uint32_t j, float& a, float& b) {
    const uint64_t z = rng_bits(key, j);
    a = rng_u01((uint32_t)(z >> 32));
    b = rng_u01((uint32_t)z);
}

// ------------------------------------------------------------------------------------------------
// warp-aggregated queue append: one atomicAdd per warp (ballot + popc + shuffle of the base).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_append(bool pred, uint32_t* counter) {
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0) return 0;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(mask & ((1u << lane) - 1u));
}

}   // namespace tpt
