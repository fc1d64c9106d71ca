// env.cuh — importance sampling of the lat-long environment map by an alias table.
//
// The reference builds a luminance*theta "pdf" and a CDF for its environment map (src/env_light.cu:10-36,
// include/env_light.cuh:21-70) but never samples them: EnvLightSample / EnvLightPDF have no call site, its theta
// factor should be sin(theta), and the row-CDF indexing is wrong.  Here the distribution is
// luminance * sin(theta) per texel (the solid angle of a lat-long texel is (2pi/W)(pi/H) sin(theta)), sampled in O(1)
// with Walker/Vose alias tables instead of two binary searches, and used for OPTIONAL next-event estimation
// (TPT_RENDER_ENV_NEE) combined with the reference's cosine-sampled paths by the balance heuristic, which leaves
// the reference's expectation (environment light arrives through path misses) unchanged.
//
// Texel <-> direction is the mapping of sampleEnvLights / Vec2UV (src/path_tracer.cu:288-294,
// include/env_light.cuh:72-78): u = atan2(d.z, d.x) / 2pi (+1 if negative), v = 1 - acos(d.y) / pi, row 0 = bottom.
#pragma once
#include "common.cuh"

namespace tpt {

struct EnvSampler {
    const float* prob;        // alias table: acceptance probability of texel i
    const uint32_t* alias;    // alias table: the other texel of bucket i
    const float* mass;        // probability mass of texel i (sums to 1)
    int w, h;
};

// Draw one texel and a point inside it.  bucketBits (32 uniform bits) picks the alias-table bucket, `coin` in (0, 1] — an
// INDEPENDENT uniform — decides between the bucket's own texel and its alias, (ux, uy) place the point inside the texel.
// (Bucket and coin used to come from one 24-bit float: with the 2^23 texels of a 4096 x 2048 map the coin was left with one
// bit, and the sampled distribution was not `mass`.)  Returns the unit direction and its density per unit solid angle.
__device__ __forceinline__ float3 env_sample_dir(const EnvSampler& es, uint32_t bucketBits, float coin, float ux, float uy, float& pdf) {
    const uint32_t n = (uint32_t)es.w * (uint32_t)es.h;
    uint32_t i = min(__umulhi(bucketBits, n), n - 1u);
    if (coin > __ldg(es.prob + i)) i = __ldg(es.alias + i);   // P(coin <= p) = floor(p * 2^24) / 2^24
    const uint32_t y = i / (uint32_t)es.w, x = i - y * (uint32_t)es.w;
    const float u = ((float)x + fminf(ux, 0.99999994f)) / (float)es.w, v = ((float)y + fminf(uy, 0.99999994f)) / (float)es.h;
    const float theta = (1.0f - v) * kPi, phi = 2.0f * kPi * u;
    float st, ct, sp, cp;
    sincosf(theta, &st, &ct);
    sincosf(phi, &sp, &cp);
    // uniform in (u, v) inside the texel: density mass * n per unit (u, v), and d(omega) = 2 pi^2 sin(theta) du dv
    pdf = __ldg(es.mass + i) * (float)n / (2.0f * kPi * kPi * fmaxf(st, 1e-6f));
    return f3(st * cp, ct, st * sp);
}

// Density (per unit solid angle) with which env_sample_dir produces direction d (any length).
__device__ __forceinline__ float env_pdf(const EnvSampler& es, float3 d) {
    const float len = sqrtf(dot(d, d));
    float u = atan2f(d.z, d.x) / (2.0f * kPi);
    if (u < 0.0f) u += 1.0f;
    const float ct = fmaxf(fminf(d.y / len, 1.0f), -1.0f);
    const float v = 1.0f - acosf(ct) / kPi;
    int x = (int)floorf(u * (float)es.w), y = (int)floorf(v * (float)es.h);
    x = min(max(x, 0), es.w - 1);
    y = min(max(y, 0), es.h - 1);
    const float st = fmaxf(sqrtf(d.x * d.x + d.z * d.z) / len, 1e-6f);
    const float n = (float)es.w * (float)es.h;
    return __ldg(es.mass + (size_t)y * es.w + x) * n / (2.0f * kPi * kPi * st);
}

}   // namespace tpt
