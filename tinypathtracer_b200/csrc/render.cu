// render.cu — wavefront path tracer for sm_100a.
//
// Replaces the `trace` megakernel of the reference (src/path_tracer.cu:296-435), its ray generator
// sampleRays (:42-59), BSDF sampling getNewDirection (:187-225), delta-light NEE sampleDeltaLights
// (:265-286), env lookup sampleEnvLights (:288-294) and copyToFB (:451-471).
//
// Kernels per batch (= all pixels x k samples):
//   k_primary   wave 0: generate + closest hit + shade of the camera vertex, fused.  Primary rays are
//               coherent (a warp = an 8x4 pixel tile), so nothing is gained by queueing them; paths
//               that miss or hit an emitter end in registers (per-thread int64 sums, one atomic per
//               pixel and batch) and never touch a queue.
//   k_wave      wave w >= 1: one persistent launch consumes the two queues the previous wave filled —
//               extension rays towards vertex w (closest hit -> shade -> emit) and the secondary rays
//               of vertex w-1 (shadow rays: any hit; the reference's extra cosine "direct" rays:
//               closest hit -> emission) — 32-item chunks handed to warps by an atomic cursor,
//               output compacted by warp-aggregated append.
//
// Estimator.  The reference unwinds  radiance = (direct[d] + radiance) * atten[d] / p[d]  backwards
// over a per-thread stack.  With T_d = prod_{k<=d} atten_k / p_k this is, forwards,
//     L = sum_d T_d * direct_d + T_{D-1} * L_end,
// where atten/p == baseColor for all three lobes (diffuse: (|c|/pi) / (c/pi) with c > 0),
// direct_d = sum_lights visible * baseColor * Li  (+ emission of whatever the extra cosine "direct"
// ray hits, diffuse vertices only), and L_end is env(dir) on a miss, the grey emission when the
// vertex is emissive (which also discards that vertex's own direct term, :421-424), or 0 at the depth
// cut.  Every term is added to the pixel as an exact fixed-point integer, so the result does not
// depend on the order in which kernels, warps or GPUs finish, nor on how terms are grouped.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>

#include <cooperative_groups.h>

#include "env.cuh"
#include "traverse.cuh"

namespace tpt {

// ---- path state: 3 x float4 per path ---------------------------------------------------------
//   s0 = (o.x, o.y, o.z, d.x)   s1 = (d.y, d.z, T.r, T.g)   s2 = (T.b, pixel, sample, -)
// secondary (shadow / direct) ray: same shape, T replaced by the contribution weight, s2.w = kind.
struct PathRec {
    float3 o, d, T;
    uint32_t pixel, sample;
    uint32_t w;   // extension ray, env NEE kernels: float bits of the solid-angle pdf its direction was drawn with, 0 = delta / camera;
                  // other kernels: how many vertices AHEAD of its wave the path is (k_primary follows specular chains in registers,
                  // so the ray consumed by wave w heads for vertex w + ahead)
};
__device__ __forceinline__ PathRec load_path(const float4* __restrict__ p, uint32_t i) {
    const float4 a = __ldcs(p + 3 * (size_t)i), b = __ldcs(p + 3 * (size_t)i + 1), c = __ldcs(p + 3 * (size_t)i + 2);
    PathRec r;
    r.o = f3(a.x, a.y, a.z);
    r.d = f3(a.w, b.x, b.y);
    r.T = f3(b.z, b.w, c.x);
    r.pixel = __float_as_uint(c.y);
    r.sample = __float_as_uint(c.z);
    r.w = __float_as_uint(c.w);
    return r;
}
__device__ __forceinline__ void store_path(float4* __restrict__ p, uint32_t i, float3 o, float3 d, float3 T, uint32_t pixel,
                                           uint32_t sample, uint32_t w) {
    __stcs(p + 3 * (size_t)i, make_float4(o.x, o.y, o.z, d.x));
    __stcs(p + 3 * (size_t)i + 1, make_float4(d.y, d.z, T.x, T.y));
    __stcs(p + 3 * (size_t)i + 2, make_float4(T.z, __uint_as_float(pixel), __uint_as_float(sample), __uint_as_float(w)));
}

// ---- accumulation: 3 x int64 fixed point per pixel ----------------------------------------------
// One contribution channel -> fixed point: min(c, 2^26) * 2^36, round to nearest.  Integer sums are
// exact, so summing in registers first and adding once gives the same bits as one atomic per term.
__device__ __forceinline__ long long to_fixed(float c) {
    const float scale = 68719476736.0f;   // 2^36
    const float cap = 67108864.0f;        // 2^26 per contribution keeps int64 far from overflow
    return __float2ll_rn(fminf(c, cap) * scale);
}
__device__ __forceinline__ bool finite3(float3 c) { return isfinite(c.x) && isfinite(c.y) && isfinite(c.z); }
__device__ __forceinline__ void accum_add(unsigned long long* __restrict__ acc, uint32_t pixel, float3 c, DeviceCounters* ctr) {
    if (!finite3(c)) {
        atomicAdd(&ctr->dropped_nonfinite, 1ull);
        return;
    }
    unsigned long long* p = acc + 3 * (size_t)pixel;
    if (c.x != 0.f) atomicAdd(p + 0, (unsigned long long)to_fixed(c.x));
    if (c.y != 0.f) atomicAdd(p + 1, (unsigned long long)to_fixed(c.y));
    if (c.z != 0.f) atomicAdd(p + 2, (unsigned long long)to_fixed(c.z));
}

// ---- queues -----------------------------------------------------------------------------------------
// Every queue is split into kSegments independent SEGMENTS, each with its own counters on its own 128-B line: warps
// of the whole GPU appending through one counter serialise in the L2 atomic unit (ncu: half of k_primary's stall
// samples were the two append atomics), spread over a few lines they do not (8 segments measured as good as 64).  A path stays in the segment its camera
// sample was born in (k_primary: blockIdx % kSegments), so segment s of any level holds at most as many extension
// rays as segment s of level 1 can — the capacity bound that makes fixed segment sizes safe.
// A segment's counter block describes what wave w consumes from it.  Items are 3 float4 (PathRec above); secondary
// items carry the contribution weight in T and their kind in s2.w.  Both item counts live in ONE 64-bit word so
// that a warp reserves the slots of everything a group of vertices emits with a single atomic round trip.
#ifndef TPT_QUEUE_SEGMENTS
#define TPT_QUEUE_SEGMENTS 8
#endif
constexpr int kSegments = TPT_QUEUE_SEGMENTS;   // power of two
constexpr int kLevelStride = 128;               // bytes between the counter blocks of two segments
struct QueueLevel {
    unsigned long long counts;       // low 32 bits: extension rays towards vertex w; high 32 bits: secondary rays of vertex w-1
    uint32_t cursor;                 // segment 0 only: next item of the whole level not yet handed to a warp (k_wave)
    uint32_t n_direct;               // how many of the secondary rays are "direct" rays (statistics only)
};
__device__ __forceinline__ QueueLevel* seg_level(void* level, uint32_t seg) {
    return reinterpret_cast<QueueLevel*>(static_cast<char*>(level) + (size_t)seg * kLevelStride);
}
constexpr uint32_t kKindShadow = 0u, kKindDirect = 1u;

// ---- pixel enumeration: 8x4 tiles so that a warp's primary rays are coherent ---------------------
__host__ __device__ __forceinline__ uint32_t tiles_x(int W) { return (uint32_t)(W + 7) >> 3; }
__host__ __device__ __forceinline__ uint32_t tiles_y(int H) { return (uint32_t)(H + 3) >> 2; }
__device__ __forceinline__ bool tile_decode(uint32_t p, int W, int H, int& x, int& y) {
    const uint32_t tile = p >> 5, lane = p & 31u, tx = tiles_x(W);
    x = (int)((tile % tx) * 8u + (lane & 7u));
    y = (int)((tile / tx) * 4u + (lane >> 3));
    return x < W && y < H;
}

// sampleRays, src/path_tracer.cu:42-59: pixel row 0 is the bottom of the image; aspect is the
// camera's, not W/H; the direction is normalised with ref_frsqrt.
struct CameraParams {
    float c2w[16];
    float sensorW, sensorH;
    float invW, invH;
};
__device__ __forceinline__ float3 primary_dir(const CameraParams& cam, int px, int py, float ju, float jv) {
    // one rounding per operation, in the reference's order (bit-identical primary rays)
    float plx = xmul(xadd(ju, (float)px), cam.invW);
    float ply = xmul(xadd(jv, (float)py), cam.invH);
    plx = xmul(plx, cam.sensorW);
    ply = xmul(ply, cam.sensorH);
    const float3 dir = f3(xsub(plx, xmul(0.5f, cam.sensorW)), xsub(ply, xmul(0.5f, cam.sensorH)), -1.0f);
    return xnormalize(xmat4_dir(cam.c2w, dir));
}
__device__ __forceinline__ void primary_ray(const CameraParams& cam, int px, int py, float ju, float jv, float3& o, float3& d) {
    d = primary_dir(cam, px, py, ju, jv);
    o = xmat4_point(cam.c2w, f3(0.f, 0.f, 0.f));
}

// The pyramid that contains every jittered ray of pixel (px, py): see traverse.cuh "pixel-frustum culling".
__device__ __forceinline__ PixelFrustum pixel_frustum(const CameraParams& cam, int px, int py, float3 o) {
    const float eps = 1.0f / 64.0f;
    const float x0 = ((float)px - eps) * cam.invW * cam.sensorW - 0.5f * cam.sensorW, x1 = ((float)px + 1.0f + eps) * cam.invW * cam.sensorW - 0.5f * cam.sensorW;
    const float y0 = ((float)py - eps) * cam.invH * cam.sensorH - 0.5f * cam.sensorH, y1 = ((float)py + 1.0f + eps) * cam.invH * cam.sensorH - 0.5f * cam.sensorH;
    float3 c[4] = {mat4_dir(cam.c2w, f3(x0, y0, -1.f)), mat4_dir(cam.c2w, f3(x1, y0, -1.f)), mat4_dir(cam.c2w, f3(x1, y1, -1.f)),
                   mat4_dir(cam.c2w, f3(x0, y1, -1.f))};
    const float3 centre = (c[0] + c[1]) + (c[2] + c[3]);
    PixelFrustum f;
    f.o = o;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float3 n = cross(c[k], c[(k + 1) & 3]);
        if (dot(n, centre) < 0.f) n = -n;   // inward, whatever the handedness of c2w
        f.n[k] = n;
    }
    return f;
}

__global__ void __launch_bounds__(256) k_primary_rays(CameraParams cam, int W, int H, int jitter, uint64_t seedKey, tpt_ray* __restrict__ rays) {
    const uint32_t n = (uint32_t)W * (uint32_t)H;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int px = (int)(i % (uint32_t)W), py = (int)(i / (uint32_t)W);
        float ju = 0.5f, jv = 0.5f;
        if (jitter) rng_pair(rng_path_key(seedKey, i, 0u), 0u, ju, jv);
        float3 o, d;
        primary_ray(cam, px, py, ju, jv, o, d);
        tpt_ray r;
        r.origin[0] = o.x; r.origin[1] = o.y; r.origin[2] = o.z;
        r.dir[0] = d.x; r.dir[1] = d.y; r.dir[2] = d.z;
        rays[i] = r;
    }
}

// ---- batch trace entry (tpt_trace_rays) -------------------------------------------------------------
template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(128) k_trace_rays(BvhView bvh, const tpt_ray* __restrict__ rays, size_t n, tpt_hit* __restrict__ hits,
                                                    DeviceCounters* ctr) {
    int stack[kStackSize];
    stack[0] = kStackSentinel;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float* r = reinterpret_cast<const float*>(rays + i);
        const float2 a = __ldg(reinterpret_cast<const float2*>(r)), b = __ldg(reinterpret_cast<const float2*>(r) + 1),
                     c = __ldg(reinterpret_cast<const float2*>(r) + 2);
        const HitRec h = traverse<COUNT>(bvh, f3(a.x, a.y, b.x), f3(b.y, c.x, c.y), ANY, stack, &ctr->node_visits, &ctr->tri_tests);
        reinterpret_cast<float4*>(hits)[i] = make_float4(__int_as_float(h.prim), h.t, h.u, h.v);
    }
}

// ---- BSDF pieces (reference src/path_tracer.cu:137-225, include/sampler.h:75-96) --------------------
__device__ __forceinline__ float3 reflect_dir(float3 d, float3 n) { return d - 2.0f * dot(d, n) * n; }

__device__ __forceinline__ float3 hemisphere_cosine(float xi1, float xi2, float3 n) {
    float3 xBase = (n.z == 0.0f) ? f3(0.f, 0.f, 1.f) : f3(1.0f, 0.0f, -n.x / n.z);
    xBase = (1.0f / sqrtf(dot(xBase, xBase))) * xBase;
    const float3 zBase = cross(xBase, n);
    const float phi = 2.0f * kPi * xi1;
    const float cosT = sqrtf(xi2);
    const float sinT = sqrtf(1.0f - cosT * cosT);
    float sp, cp;
    sincosf(phi, &sp, &cp);
    return (cp * sinT) * xBase + cosT * n + (sp * sinT) * zBase;
}

// Rough metal lobe (TPT_RENDER_ROUGH_METALS; not in the reference, whose metals are mirrors): GGX with alpha = roughness^2,
// half-vector drawn from the distribution of VISIBLE normals (Heitz 2018) around the shading normal facing the ray, reflection
// about it.  With Fresnel = baseColor (as for the mirror) the sample weight is baseColor * G1(wi), G1 the Smith masking term;
// a direction below the surface has weight 0.  Returns G1(wi).  Local frame = the one hemisphere_cosine builds.
__device__ __forceinline__ float rough_metal_direction(float3 rayDir, float3 n, float alpha, float xi1, float xi2, float3& out) {
    const float3 nn = dot(rayDir, n) > 0.0f ? -n : n;
    float3 xBase = (nn.z == 0.0f) ? f3(0.f, 0.f, 1.f) : f3(1.0f, 0.0f, -nn.x / nn.z);
    xBase = (1.0f / sqrtf(dot(xBase, xBase))) * xBase;
    const float3 zBase = cross(xBase, nn);
    const float3 wo = f3(-dot(rayDir, xBase), -dot(rayDir, zBase), -dot(rayDir, nn));   // towards the viewer, local (x, y, z = normal)
    float3 vh = f3(alpha * wo.x, alpha * wo.y, wo.z);
    vh = (1.0f / sqrtf(dot(vh, vh))) * vh;
    const float lensq = vh.x * vh.x + vh.y * vh.y;
    const float3 t1 = lensq > 0.0f ? (1.0f / sqrtf(lensq)) * f3(-vh.y, vh.x, 0.0f) : f3(1.0f, 0.0f, 0.0f);
    const float3 t2 = cross(vh, t1);
    const float r = sqrtf(xi1), phi = 2.0f * kPi * xi2;
    float sphi, cphi;
    sincosf(phi, &sphi, &cphi);
    const float p1 = r * cphi, sh = 0.5f * (1.0f + vh.z);
    const float p2 = (1.0f - sh) * sqrtf(fmaxf(1.0f - p1 * p1, 0.0f)) + sh * (r * sphi);
    const float3 nh = p1 * t1 + p2 * t2 + sqrtf(fmaxf(1.0f - p1 * p1 - p2 * p2, 0.0f)) * vh;
    float3 ne = f3(alpha * nh.x, alpha * nh.y, fmaxf(nh.z, 0.0f));
    ne = (1.0f / sqrtf(dot(ne, ne))) * ne;
    const float3 wi = 2.0f * dot(wo, ne) * ne - wo;
    out = wi.x * xBase + wi.y * zBase + wi.z * nn;
    if (!(wi.z > 0.0f)) return 0.0f;
    return 2.0f * wi.z / (wi.z + sqrtf(alpha * alpha + (1.0f - alpha * alpha) * wi.z * wi.z));
}

// getNewDirection, src/path_tracer.cu:187-225.  Returns false when the reference's pdf would be 0
// (cosine sample below the horizon after rounding: its result is Inf/NaN; we drop the sub-path).
__device__ __forceinline__ bool new_direction(float3 rayDir, float3 n, float eta_m, float metallic, float xi1, float xi2, float3& out) {
    if (eta_m > 0.0f) {
        float cosI = dot(rayDir, n);
        const float eta = cosI > 0.0f ? eta_m : 1.0f / eta_m;
        const float3 nn = cosI > 0.0f ? -n : n;
        const float3 refl = reflect_dir(rayDir, n);
        cosI = fabsf(cosI);
        const float sin2T = eta * eta * (1.0f - cosI * cosI);
        float fr = 1.0f;
        float3 refr = f3(0.f, 0.f, 0.f);
        if (sin2T < 1.0f) {
            const float cosT = sqrtf(1.0f - sin2T);
            refr = eta * rayDir + (cosI * eta - cosT) * nn;
            float f0 = (1.0f - eta) / (1.0f + eta);
            f0 *= f0;
            const float m = fmaxf(fminf(1.0f - cosI, 1.0f), 0.0f);
            const float m2 = m * m;
            fr = f0 + (1.0f - f0) * m2 * m2 * m;
        }
        out = (xi1 < fr) ? refl : refr;
        return true;
    }
    if (metallic > 0.0f) {
        out = reflect_dir(rayDir, n);
        return true;
    }
    const float3 nn = dot(rayDir, n) > 0.0f ? -n : n;
    out = hemisphere_cosine(xi1, xi2, nn);
    return dot(out, nn) > 0.0f;
}

// DeltaLight::sample + CalcDistAttenuation, include/delta_light.h:25-127
__device__ __forceinline__ void light_incoming(const tpt_delta_light& L, float3 p, float3& dir, float3& radiance) {
    const float3 color = f3(L.color[0], L.color[1], L.color[2]);
    float dist = 0.0f;
    if (L.type == 1) {
        dir = -f3(L.a[0], L.a[1], L.a[2]);
        radiance = color * L.intensity;
    } else {
        dir = f3(L.a[0], L.a[1], L.a[2]) - p;
        dist = sqrtf(dot(dir, dir));
        dir = (1.0f / dist) * dir;
        radiance = color * L.intensity;
        if (L.type == 2) {
            const float cosT = dot(-dir, f3(L.b[0], L.b[1], L.b[2]));
            const float s = fmaxf(fminf(cosT - L.cos_outer, 1.0f), 0.0f) * L.inv_cone_diff;
            radiance = radiance * (s * s);
        }
    }
    const float d2 = dist * dist;
    float att = 1.0f / (d2 + 1.0f);
    const float w = d2 * 0.01f;
    const float sat = fmaxf(fminf(1.0f - w * w, 1.0f), 0.0f);
    att *= sat * sat;
    radiance = radiance * att;
}

// sampleEnvLights + Vec2UV, src/path_tracer.cu:288-294, include/env_light.cuh:72-78; point filter,
// clamp addressing, normalised coordinates (src/texture.cu:160-168).  Two texel formats: the reference's RGBA8
// (value / 255) and linear float RGB for HDR maps (tpt_env_set_rgb32f; one 128-bit load per lookup).
// The two inverse trigonometric functions only choose a texel, so they are polynomials here (|error| < 3e-7 rad, i.e. < 2e-4 of a
// texel of a 4096 x 2048 map; libm's atan2f + acosf were 11 % of k_primary's instructions on ball.gltf):
//   atan(t), t in [0,1]: t * P8(t^2), Chebyshev fit, max error 1.1e-7;  acos(x), x in [0,1]: sqrt(1-x) * P7(x) (Abramowitz & Stegun
//   4.4.46), max error 2.5e-7 in fp32.
__device__ __forceinline__ float atan2_turns(float y, float x) {   // atan2(y, x) / 2pi in [0, 1)
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float t = mx > 0.0f ? __fdividef(mn, mx) : 0.0f;
    const float z = t * t;
    float p = 0.0028340641874819994f;
    p = fmaf(p, z, -0.016005029901862144f);
    p = fmaf(p, z, 0.042587608098983765f);
    p = fmaf(p, z, -0.07495445758104324f);
    p = fmaf(p, z, 0.10636754333972931f);
    p = fmaf(p, z, -0.14202570915222168f);
    p = fmaf(p, z, 0.19992484152317047f);
    p = fmaf(p, z, -0.3333306610584259f);
    p = fmaf(p, z, 1.0f);
    float a = p * t;                            // [0, pi/4]
    if (ay > ax) a = 0.5f * kPi - a;            // [0, pi/2]
    if (x < 0.0f) a = kPi - a;                  // [0, pi]
    const float u = a * (0.5f / kPi);
    return y < 0.0f ? 1.0f - u : u;             // atan2 < 0 -> + 1 turn
}
__device__ __forceinline__ float acos_fast(float x) {   // x in [-1, 1]
    const float ax = fabsf(x);
    float p = -0.0012624911f;
    p = fmaf(p, ax, 0.0066700901f);
    p = fmaf(p, ax, -0.0170881256f);
    p = fmaf(p, ax, 0.0308918810f);
    p = fmaf(p, ax, -0.0501743046f);
    p = fmaf(p, ax, 0.0889789874f);
    p = fmaf(p, ax, -0.2145988016f);
    p = fmaf(p, ax, 1.5707963050f);
    const float r = sqrtf(1.0f - ax) * p;
    return x < 0.0f ? kPi - r : r;
}
__device__ __forceinline__ float3 env_lookup(const uchar4* __restrict__ env, const float4* __restrict__ envf, int ew, int eh, float3 d) {
    if (env == nullptr && envf == nullptr) return f3(0.f, 0.f, 0.f);
    const float u = atan2_turns(d.z, d.x);
    const float v = 1.0f - acos_fast(fmaxf(fminf(d.y, 1.0f), -1.0f)) * (1.0f / kPi);
    int x = (int)floorf(u * (float)ew), y = (int)floorf(v * (float)eh);
    x = min(max(x, 0), ew - 1);
    y = min(max(y, 0), eh - 1);
    if (envf != nullptr) {
        const float4 t = __ldg(envf + (size_t)y * ew + x);
        return f3(t.x, t.y, t.z);
    }
    const uchar4 t = __ldg(env + (size_t)y * ew + x);
    return f3((float)t.x, (float)t.y, (float)t.z) * (1.0f / 255.0f);
}

struct ShadeParams {
    const float4* mtl;       // 2 float4 per material
    const tpt_delta_light* lights;
    int nLights;
    const uchar4* env;       // RGBA8 map (the reference's format) or
    const float4* envf;      // float RGB map (HDR); at most one of the two is set
    int envW, envH;
    __device__ __forceinline__ bool has_env() const { return env != nullptr || envf != nullptr; }
    int depth, maxDepth;
    uint64_t seedKey;
    int directRays;          // 0 when no material of the scene is emissive: the extra cosine "direct" ray can only add emission
    EnvSampler es;           // alias tables of the environment map (TPT_RENDER_ENV_NEE only)
    int continueSpecular;    // how many mirror / glass bounces after the camera vertex k_primary<SPEC> follows in registers
    int continueVote;        // ... when at least continueVote / 4 of the lanes that are there together want to
    int roughMetals;         // TPT_RENDER_ROUGH_METALS: metals with roughness^2 >= 1e-3 scatter by GGX instead of mirroring
    int pixelCandidates;     // k_primary: per-pixel candidate triangles instead of one tree traversal per camera sample
    int noTileWalk;          // tuning switch (TPT_NO_TILE_WALK): every pixel walks the tree itself instead of filtering its tile's leaf list
};

// ---- one path vertex ------------------------------------------------------------------------------
// Shades the vertex an extension ray (o, d, throughput T) arrives at (steps 1-7 of the reference's bounce
// loop, src/path_tracer.cu:353-414).  Returns false when the path ends here, with `terminal` = the
// contribution to add (env on a miss, grey emission on an emitter); returns true with the data needed to
// emit the next extension ray, one shadow ray per delta light and the extra cosine "direct" ray.
struct Vertex {
    float3 P;        // hit point
    float3 Tn;       // throughput after this vertex = T * baseColor
    float3 base;     // baseColor (delta-light terms carry it twice, SURVEY §3.3 iii)
    float3 ndir;     // next direction
    float3 ddir;     // direction of the "direct" ray
    bool next, direct;
    bool delta;      // ndir came from a mirror / glass lobe: one determined direction, nothing to stage or sort
    // TPT_RENDER_ENV_NEE only: pdf of ndir (0 = delta lobe), and the environment sample (direction, weighted contribution)
    float pb;
    bool nee;
    float3 ldir, lw;
};

//
// NEE (TPT_RENDER_ENV_NEE): the environment term of the estimator, E[Tn * env(ndir) * miss], is evaluated by two
// strategies combined with the balance heuristic — the reference's cosine-sampled extension ray (weight
// pb / (pb + pl) applied where it misses) and one alias-table sample of the map per diffuse vertex (a shadow ray of
// weight Tn * env(l) * pb(l) / (pl(l) + pb(l)), pb(l) = cos/pi).  Delta lobes and camera rays keep weight 1.
// The expectation is the reference's; only the variance changes.
template <bool NEE>
__device__ __forceinline__ bool shade_vertex(const BvhView& bvh, const ShadeParams& sp, int depth, float3 o, float3 d, float3 T, uint32_t pbBits,
                                             uint32_t pixel, uint32_t sample, const HitRec& h, float3& terminal, Vertex& vx,
                                             DeviceCounters* ctr) {
    if (h.leaf < 0) {
        terminal = T * env_lookup(sp.env, sp.envf, sp.envW, sp.envH, d);
        if (NEE && pbBits != 0u) {
            const float pb = __uint_as_float(pbBits);
            terminal = terminal * (pb / (pb + env_pdf(sp.es, d)));
        }
        return false;
    }
    const float4 g2 = __ldg(bvh.tri_geom + 3 * (size_t)h.leaf + 2);
    const int mtlIdx = __float_as_int(g2.z);
    const float4 m0 = __ldg(sp.mtl + 2 * mtlIdx), m1 = __ldg(sp.mtl + 2 * mtlIdx + 1);
    if (m0.w > 0.0f) {   // emissive vertex: terminal, grey emission
        terminal = T * m0.w;
        return false;
    }
    terminal = f3(0.f, 0.f, 0.f);
    const float4 na = __ldg(bvh.tri_norm + 3 * (size_t)h.leaf), nb = __ldg(bvh.tri_norm + 3 * (size_t)h.leaf + 1),
                 nc = __ldg(bvh.tri_norm + 3 * (size_t)h.leaf + 2);
    const float u = h.u, v = h.v, w = xsub(xsub(1.0f, u), v);
    const float3 n0 = f3(na.x, na.y, na.z), n1 = f3(na.w, nb.x, nb.y), n2 = f3(nb.z, nb.w, nc.x);
    const float3 nrm = xnormalize(f3(xadd(xadd(xmul(w, n0.x), xmul(u, n1.x)), xmul(v, n2.x)),
                                     xadd(xadd(xmul(w, n0.y), xmul(u, n1.y)), xmul(v, n2.y)),
                                     xadd(xadd(xmul(w, n0.z), xmul(u, n1.z)), xmul(v, n2.z))));
    vx.P = f3(xadd(o.x, xmul(h.t, d.x)), xadd(o.y, xmul(h.t, d.y)), xadd(o.z, xmul(h.t, d.z)));
    vx.base = f3(m0.x, m0.y, m0.z);
    vx.Tn = T * vx.base;
    const float eta = m1.x, metallic = m1.y;
    const uint64_t key = rng_path_key(sp.seedKey, pixel, sample);
    vx.direct = sp.directRays && !(eta >= 1.0f || metallic > 0.0f);
    vx.next = depth + 1 < sp.maxDepth;
    vx.ddir = vx.ndir = f3(0.f, 0.f, 1.f);
    if (vx.direct) {   // the reference's extra cosine "direct" ray (diffuse only)
        float a, b;
        rng_pair(key, 2u + 2u * (uint32_t)depth, a, b);
        vx.direct = new_direction(d, nrm, eta, metallic, a, b, vx.ddir);
        if (!vx.direct) atomicAdd(&ctr->dropped_nonfinite, 1ull);
    }
    const float alpha = m1.z * m1.z;
    vx.delta = eta > 0.0f || metallic > 0.0f;
    if (sp.roughMetals && metallic > 0.0f && !(eta > 0.0f) && alpha >= 1e-3f) {
        vx.delta = false;
        // drawn at the last vertex too: G1 is the estimator's attenuation at this vertex and scales its light terms as well,
        // as every attenuation does in the reference's unwind (src/path_tracer.cu:417-430)
        float a, b;
        rng_pair(key, 1u + 2u * (uint32_t)depth, a, b);
        vx.Tn = vx.Tn * rough_metal_direction(d, nrm, alpha, a, b, vx.ndir);
        vx.next = vx.next && (vx.Tn.x > 0.f || vx.Tn.y > 0.f || vx.Tn.z > 0.f);
    } else if (vx.next) {
        float a, b;
        rng_pair(key, 1u + 2u * (uint32_t)depth, a, b);
        vx.next = new_direction(d, nrm, eta, metallic, a, b, vx.ndir);
        if (!vx.next) atomicAdd(&ctr->dropped_nonfinite, 1ull);
        vx.next = vx.next && (vx.Tn.x > 0.f || vx.Tn.y > 0.f || vx.Tn.z > 0.f);   // a black path adds nothing further
    }
    if (NEE) {
        vx.pb = 0.0f;
        vx.nee = false;
        if (vx.next && !(eta > 0.0f) && !(metallic > 0.0f) && sp.has_env()) {   // diffuse lobe of new_direction
            const float3 nn = dot(d, nrm) > 0.0f ? -nrm : nrm;
            vx.pb = dot(vx.ndir, nn) * (1.0f / kPi);
            float u1, u2, u3, u4, pl;
            rng_pair(key, 0x10000u + 2u * (uint32_t)depth, u1, u2);
            rng_pair(key, 0x10001u + 2u * (uint32_t)depth, u3, u4);
            vx.ldir = env_sample_dir(sp.es, (uint32_t)(rng_bits(key, 0x10000u + 2u * (uint32_t)depth) >> 32), u4, u2, u3, pl);
            const float pbl = dot(vx.ldir, nn) * (1.0f / kPi);
            if (pbl > 0.0f && pl > 0.0f) {
                vx.lw = (vx.Tn * env_lookup(sp.env, sp.envf, sp.envW, sp.envH, vx.ldir)) * (pbl / (pl + pbl));
                vx.nee = vx.lw.x > 0.f || vx.lw.y > 0.f || vx.lw.z > 0.f;
            }
        }
    }
    return true;
}

struct QueueOut {
    float4* ext;
    float4* sec;
    void* level;                       // kSegments counter blocks of the level being produced
    uint32_t extSegCap, secSegCap;     // items per segment
    uint32_t extCap, secCap;           // items the producing kernel may write per segment it uses: the segment for k_primary, the whole
                                       // pool (k_wave writes segment 0 only and ignores the segment borders)
};

// Emit the rays of a shaded vertex from whatever lanes are here (diverged callers allowed): the lanes of one
// segment that are here together reserve all their slots — one extension ray, one shadow ray per delta light, the
// "direct" ray, the environment sample — with one atomic.  UNIFORM: every lane of the warp is in the same segment.
template <bool NEE, bool UNIFORM, bool LIGHTS = true>
__device__ __forceinline__ void emit_vertex(const ShadeParams& sp, const QueueOut& q, uint32_t seg, const Vertex& vx, uint32_t pixel,
                                            uint32_t sample, uint32_t ahead, DeviceCounters* ctr) {
    const unsigned here = __activemask();
    const unsigned m = UNIFORM ? here : __match_any_sync(here, seg);
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned bExt = __ballot_sync(here, vx.next) & m, bDir = __ballot_sync(here, vx.direct) & m,
                   bNee = NEE ? (__ballot_sync(here, vx.nee) & m) : 0u;
    const uint32_t nL = LIGHTS ? (uint32_t)sp.nLights : 0u;   // !LIGHTS: the caller has traced the vertex's light shadow rays itself
    const uint32_t nExt = __popc(bExt), nSec = nL * __popc(m) + __popc(bDir) + __popc(bNee);
    unsigned long long base = 0;
    if (lane == leader && nExt + nSec != 0u) {
        QueueLevel* level = seg_level(q.level, seg);
        base = atomicAdd(&level->counts, (unsigned long long)nExt | ((unsigned long long)nSec << 32));
        if (bDir) atomicAdd(&level->n_direct, (uint32_t)__popc(bDir));   // result unused: a fire-and-forget reduction
    }
    base = __shfl_sync(m, base, leader);
    float4* ext = q.ext + 3 * (size_t)seg * q.extSegCap;
    float4* sec = q.sec + 3 * (size_t)seg * q.secSegCap;
    uint32_t slot = (uint32_t)(base >> 32) + nL * __popc(m & lt) + __popc(bDir & lt) + __popc(bNee & lt);
    // The pools are sized for the most every path can emit, so a reservation cannot pass the end; should that ever be wrong,
    // nothing is written out of bounds and nothing is left unwritten below the cap: every reserved slot that lies inside its pool
    // is written (the consumer clamps its counts to the cap, so it reads exactly those), the rest is dropped and counted.
    const uint32_t extSlot = (uint32_t)base + __popc(bExt & lt);
    const uint32_t mySec = nL + (vx.direct ? 1u : 0u) + ((NEE && vx.nee) ? 1u : 0u);
    const bool extFits = extSlot < q.extCap;
    if (slot + mySec > q.secCap || (vx.next && !extFits)) atomicAdd(&ctr->queue_overflow, 1ull);
    // delta lights: one infinite shadow ray per light, contribution T_d * baseColor * Li
    for (int l = 0; l < (int)nL; ++l, ++slot) {
        if (slot >= q.secCap) continue;
        float3 ldir, lrad;
        light_incoming(sp.lights[l], vx.P, ldir, lrad);
        store_path(sec, slot, vx.P, ldir, vx.Tn * (vx.base * lrad), pixel, sample, kKindShadow);
    }
    if (vx.direct) {
        if (slot < q.secCap) store_path(sec, slot, vx.P, vx.ddir, vx.Tn, pixel, sample, kKindDirect);
        ++slot;
    }
    if (NEE && vx.nee) {
        if (slot < q.secCap) store_path(sec, slot, vx.P, vx.ldir, vx.lw, pixel, sample, kKindShadow);
        ++slot;
    }
    if (vx.next && extFits) store_path(ext, extSlot, vx.P, vx.ndir, vx.Tn, pixel, sample, NEE ? __float_as_uint(vx.pb) : ahead);
}

// The delta-light shadow rays of one vertex, traced where the vertex was shaded (k_primary): the lanes of a pixel group sit at
// neighbouring points and aim at the same light, so the while-while any-hit walk stays coherent, and the rays never cost a
// queue item, a second read and k_wave's scheduling.  Adds exactly what k_wave would add for them.
template <bool COUNT>
__device__ __forceinline__ void light_shadows_here(const BvhView& bvh, const ShadeParams& sp, const Vertex& vx, int* stack, long long& aR, long long& aG,
                                                   long long& aB, unsigned& nShadow, DeviceCounters* ctr) {
    for (int l = 0; l < sp.nLights; ++l) {
        float3 ldir, lrad;
        light_incoming(sp.lights[l], vx.P, ldir, lrad);
        const HitRec sh = traverse<COUNT>(bvh, vx.P, ldir, true, stack, &ctr->node_visits, &ctr->tri_tests);
        ++nShadow;
        if (sh.leaf < 0) {
            const float3 c = vx.Tn * (vx.base * lrad);
            if (finite3(c)) {
                if (c.x != 0.f) aR += to_fixed(c.x);
                if (c.y != 0.f) aG += to_fixed(c.y);
                if (c.z != 0.f) aB += to_fixed(c.z);
            } else {
                atomicAdd(&ctr->dropped_nonfinite, 1ull);
            }
        }
    }
}

// ---- wave 0: generate + extend + shade, fused ---------------------------------------------------------
// One camera sample: jitter, direction, closest hit among the pixel's candidates (cand/bound: nc entries sorted by a
// lower bound of t) or by a tree walk (nc < 0) or none at all (nc == 0), then the first vertex.  Terminal
// contributions are summed into (aR, aG, aB) in exact fixed point.
// MODE bit 0 (SPEC): specular continuation; bit 1 (LHERE): the camera-side vertices' light shadow rays are traced here.
template <bool COUNT, bool NEE, int MODE>
__device__ __forceinline__ void primary_sample(const BvhView& bvh, const ShadeParams& sp, const CameraParams& cam, const QueueOut& q, uint32_t seg,
                                               int px, int py, uint32_t pixel, uint32_t sIdx, float3 o, int nc, const int* cand, const int* bound, int candStride,
                                               int* stack, long long& aR, long long& aG, long long& aB, unsigned& nt, unsigned& nExt, unsigned& nShadow,
                                               DeviceCounters* ctr) {
    constexpr bool SPEC = (MODE & 1) != 0, LHERE = (MODE & 2) != 0;
    float ju, jv;
    rng_pair(rng_path_key(sp.seedKey, pixel, sIdx), 0u, ju, jv);
    const float3 d = primary_dir(cam, px, py, ju, jv);
    HitRec h;
    h.leaf = -1;
    if (nc > 0) {
        h.t = FLT_MAX;
        h.u = h.v = 0.0f;
        h.prim = -1;
        for (int k = 0; k < nc; ++k) {   // sorted by a lower bound of t: nothing beyond the best hit needs testing
            if (h.t < __int_as_float(bound[k * candStride])) break;
            leaf_test(bvh, cand[k * candStride], o, d, h);
            if (COUNT) ++nt;
        }
    } else if (nc < 0) {
        h = traverse<COUNT>(bvh, o, d, false, stack, &ctr->node_visits, &ctr->tri_tests);
    }
    if (!SPEC) {
        float3 term;
        Vertex vx;
        if (shade_vertex<NEE>(bvh, sp, 0, o, d, f3(1.f, 1.f, 1.f), 0u, pixel, sIdx, h, term, vx, ctr)) {
            if (LHERE) light_shadows_here<COUNT>(bvh, sp, vx, stack, aR, aG, aB, nShadow, ctr);
            emit_vertex<NEE, true, !LHERE>(sp, q, seg, vx, pixel, sIdx, 0u, ctr);
        } else if (term.x != 0.f || term.y != 0.f || term.z != 0.f) {
            if (finite3(term)) {
                aR += to_fixed(term.x);
                aG += to_fixed(term.y);
                aB += to_fixed(term.z);
            } else {
                atomicAdd(&ctr->dropped_nonfinite, 1ull);
            }
        }
        return;
    }
    // SPEC (scenes with mirror / glass materials; not the env-NEE kernels): specular continuation.  A path that leaves a vertex
    // through a mirror / glass lobe is one determined ray; when (nearly) all lanes that are here together want to go on that
    // way — a warp whose pixels lie on a mirror — the ray is traced right here with the while-while traversal instead of going
    // through a queue and the next wave, for at most sp.continueSpecular bounces (further down a chain the lanes have drifted
    // apart and k_wave's self-scheduling warps do better).  The first vertex not passed this way is queued with the number of
    // vertices it is AHEAD of its wave; terminal contributions stay in registers.  Delta-light shadow rays of the vertices
    // passed are traced here too, so a camera sample still emits the rays of at most one vertex (the pool bound).
    float3 ro = o, rd = d, T = f3(1.f, 1.f, 1.f);
    int depth = 0;
    while (true) {
        float3 term;
        Vertex vx;
        if (!shade_vertex<NEE>(bvh, sp, depth, ro, rd, T, 0u, pixel, sIdx, h, term, vx, ctr)) {
            if (term.x != 0.f || term.y != 0.f || term.z != 0.f) {
                if (finite3(term)) {
                    aR += to_fixed(term.x);
                    aG += to_fixed(term.y);
                    aB += to_fixed(term.z);
                } else {
                    atomicAdd(&ctr->dropped_nonfinite, 1ull);
                }
            }
            break;
        }
        const bool want = vx.delta && vx.next && depth < sp.continueSpecular;
        const unsigned here = __activemask();
        const unsigned wants = __ballot_sync(here, want);
        if (!want || 4 * __popc(wants) < sp.continueVote * __popc(here)) {   // a heuristic only: either way gives the same sums
            if (LHERE) light_shadows_here<COUNT>(bvh, sp, vx, stack, aR, aG, aB, nShadow, ctr);
            emit_vertex<NEE, true, !LHERE>(sp, q, seg, vx, pixel, sIdx, (uint32_t)depth, ctr);
            break;
        }
        light_shadows_here<COUNT>(bvh, sp, vx, stack, aR, aG, aB, nShadow, ctr);   // what wave depth+1 would do with this vertex's shadow rays
        ro = vx.P;
        rd = vx.ndir;
        T = vx.Tn;
        ++depth;
        ++nExt;
        h = traverse<COUNT>(bvh, ro, rd, false, stack, &ctr->node_visits, &ctr->tri_tests);
    }
}

// One walk of the tree with the pixel's pyramid finds the candidate triangles of ALL its samples (traverse.cuh):
// 0: no sample can hit anything; 1..kMaxCandidates: test those per sample; -1: too many, traverse per sample.
__device__ __forceinline__ int pixel_candidates(const BvhView& bvh, const ShadeParams& sp, const CameraParams& cam, int px, int py, float3 o,
                                                int* stack, unsigned& nv) {
    return sp.pixelCandidates ? frustum_collect(bvh, pixel_frustum(cam, px, py, o), stack, nv)
                              : (frustum_may_hit(bvh, pixel_frustum(cam, px, py, o), stack) ? -1 : 0);
}

// The pyramid of a whole 8x4 pixel tile, grown by 1/32 pixel: it contains the (1/64-pixel grown) pyramid of each of its pixels
// with 1/64 pixel to spare, so a box that passes a pixel's plane tests passes the tile's whatever the rounding (~1e-7 pixel).
__device__ __forceinline__ PixelFrustum tile_frustum(const CameraParams& cam, int tx0, int ty0, float3 o) {
    const float eps = 1.0f / 32.0f;
    const float x0 = ((float)tx0 - eps) * cam.invW * cam.sensorW - 0.5f * cam.sensorW, x1 = ((float)(tx0 + 8) + eps) * cam.invW * cam.sensorW - 0.5f * cam.sensorW;
    const float y0 = ((float)ty0 - eps) * cam.invH * cam.sensorH - 0.5f * cam.sensorH, y1 = ((float)(ty0 + 4) + eps) * cam.invH * cam.sensorH - 0.5f * cam.sensorH;
    float3 c[4] = {mat4_dir(cam.c2w, f3(x0, y0, -1.f)), mat4_dir(cam.c2w, f3(x1, y0, -1.f)), mat4_dir(cam.c2w, f3(x1, y1, -1.f)),
                   mat4_dir(cam.c2w, f3(x0, y1, -1.f))};
    const float3 centre = (c[0] + c[1]) + (c[2] + c[3]);
    PixelFrustum f;
    f.o = o;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float3 n = cross(c[k], c[(k + 1) & 3]);
        if (dot(n, centre) < 0.f) n = -n;
        f.n[k] = n;
    }
    return f;
}

// Warp-cooperative walk of the tree with the TILE's pyramid, breadth first: the 32 lanes test 32 nodes of the current frontier
// at a time, surviving internal children form the next frontier, surviving leaves are appended to the tile's leaf list (the
// parent node and side, where the leaf's box is stored, and the leaf index).  One such walk replaces 32 divergent per-pixel
// walks: each pixel then only filters the tile's (short) list with its own pyramid.  A triangle some ray of a pixel can hit
// has its box inside that pixel's pyramid, hence inside the tile's (grown by 1/64 pixel more, five orders of magnitude
// above the rounding of the plane tests), so per-pixel candidates taken from the tile list are the ones a per-pixel walk
// collects and hold every triangle a full traversal could return: the hits stay bit-identical.  Returns the number of leaves, or -1 when the frontier or the list
// overflows (the caller then walks per pixel, as before).  Warp-uniform; all 32 lanes must call it.
constexpr int kTileFront = 64, kTileLeaves = 64;
struct TileWalk {
    int front[2][kTileFront];
    int leafNode[kTileLeaves];   // (parent node << 1) | side
    int leafId[kTileLeaves];     // sorted triangle index
};
__device__ __forceinline__ int tile_collect(const BvhView& bvh, const PixelFrustum& f, TileWalk& tw, unsigned& nodeVisits) {
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    if (bvh.n_faces <= 1) {
        if (lane == 0) {
            tw.leafNode[0] = -1;   // no parent: the single triangle is always a candidate
            tw.leafId[0] = 0;
        }
        __syncwarp();
        return 1;
    }
    int cur = 0, nCur = 1, nLeaf = 0;
    if (lane == 0) tw.front[0][0] = 0;
    __syncwarp();
    while (nCur > 0) {
        int nNext = 0;
        for (int base = 0; base < nCur; base += 32) {
            const int idx = base + (int)lane;
            const bool have = idx < nCur;
            bool hL = false, hR = false;
            int2 ch = make_int2(0, 0);
            int node = 0;
            if (have) {
                node = tw.front[cur][idx];
                const float4* n = bvh.nodes + 4 * (size_t)node;
                const float4 n0 = __ldg(n), n1 = __ldg(n + 1), n2 = __ldg(n + 2);
                ch = __ldg(reinterpret_cast<const int2*>(n + 3));
                hL = frustum_box(f, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y);
                hR = frustum_box(f, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w);
                ++nodeVisits;
            }
            const bool iL = hL && ch.x >= 0, iR = hR && ch.y >= 0, lL = hL && ch.x < 0, lR = hR && ch.y < 0;
            const unsigned mIL = __ballot_sync(0xffffffffu, iL), mIR = __ballot_sync(0xffffffffu, iR);
            const unsigned mLL = __ballot_sync(0xffffffffu, lL), mLR = __ballot_sync(0xffffffffu, lR);
            const int pIL = nNext + __popc(mIL & lt), pIR = nNext + __popc(mIL) + __popc(mIR & lt);
            const int pLL = nLeaf + __popc(mLL & lt), pLR = nLeaf + __popc(mLL) + __popc(mLR & lt);
            nNext += __popc(mIL) + __popc(mIR);
            nLeaf += __popc(mLL) + __popc(mLR);
            if (nNext > kTileFront || nLeaf > kTileLeaves) return -1;   // uniform
            if (iL) tw.front[cur ^ 1][pIL] = ch.x;
            if (iR) tw.front[cur ^ 1][pIR] = ch.y;
            if (lL) { tw.leafNode[pLL] = node << 1; tw.leafId[pLL] = ~ch.x; }
            if (lR) { tw.leafNode[pLR] = (node << 1) | 1; tw.leafId[pLR] = ~ch.y; }
        }
        __syncwarp();
        cur ^= 1;
        nCur = nNext;
    }
    return nLeaf;
}

// A pixel's candidates out of its tile's leaf list: the same box test, bound and sorted insertion as frustum_collect.
// Returns like frustum_collect: the number of candidates (at the top of `stack`), -1 when there are more than kMaxCandidates.
__device__ __forceinline__ int candidates_from_tile(const BvhView& bvh, const PixelFrustum& f, const TileWalk& tw, int nLeaf, int* stack) {
    int nc = 0;
    for (int j = 0; j < nLeaf; ++j) {
        const int code = tw.leafNode[j];
        float lox, hix, loy, hiy, loz, hiz;
        if (code < 0) {   // single-triangle scene
            stack[kStackSize - 1] = 0;
            stack[kStackSize - 1 - kMaxCandidates] = 0;
            return 1;
        }
        const float4* n = bvh.nodes + 4 * (size_t)(code >> 1);
        const float4 xy = __ldg(n + (code & 1)), z = __ldg(n + 2);
        lox = xy.x; hix = xy.y; loy = xy.z; hiy = xy.w;
        loz = (code & 1) ? z.z : z.x;
        hiz = (code & 1) ? z.w : z.y;
        if (frustum_box(f, lox, hix, loy, hiy, loz, hiz))
            if (!candidate_insert(stack, nc, tw.leafId[j], box_distance_bound(f.o, lox, hix, loy, hiy, loz, hiz))) return -1;
    }
    return nc;
}

// rays k_primary traced in registers (specular continuation) never see a queue, where k_wave counts its rays
__device__ __forceinline__ void flush_ray_counters(unsigned nExt, unsigned nShadow, DeviceCounters* ctr) {
    const unsigned m = __activemask();
    const unsigned se = __reduce_add_sync(m, nExt), ss = __reduce_add_sync(m, nShadow);
    if ((threadIdx.x & 31) == __ffs(m) - 1) {
        if (se) atomicAdd(&ctr->rays_extension, (unsigned long long)se);
        if (ss) atomicAdd(&ctr->rays_shadow, (unsigned long long)ss);
    }
}

__device__ __forceinline__ void flush_trav_counters(unsigned nv, unsigned nt, DeviceCounters* ctr) {
    const unsigned m = __activemask();
    const unsigned sv = __reduce_add_sync(m, nv), st = __reduce_add_sync(m, nt);
    if ((threadIdx.x & 31) == __ffs(m) - 1) {
        atomicAdd(&ctr->node_visits, (unsigned long long)sv);
        atomicAdd(&ctr->tri_tests, (unsigned long long)st);
    }
}

// Form A (fewer than 32 samples per pixel in this launch): thread = (pixel of an 8x4 tile, group of `spt`
// consecutive local samples); the candidates live in a per-thread array of their own.
template <bool COUNT, bool NEE, int MODE>
__global__ void __launch_bounds__(128) k_primary(BvhView bvh, ShadeParams sp, CameraParams cam, int W, int H, uint32_t nPadded,
                                                 uint32_t kSamples, uint32_t spt, uint32_t sampleBase, uint32_t sampleStride, QueueOut q,
                                                 unsigned long long* __restrict__ acc, DeviceCounters* ctr) {
    const uint32_t groups = (kSamples + spt - 1) / spt;
    const uint32_t total = nPadded * groups;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    if (i == 0) atomicAdd(&ctr->rays_primary, (unsigned long long)W * (unsigned long long)H * kSamples);
    const uint32_t g = i / nPadded, p = i - g * nPadded;
    int px, py;
    if (!tile_decode(p, W, H, px, py)) return;
    const uint32_t pixel = (uint32_t)(px + py * W);
    const uint32_t s0 = g * spt, s1 = min(kSamples, s0 + spt);
    long long aR = 0, aG = 0, aB = 0;
    int stack[kStackSize];
    stack[0] = kStackSentinel;
    const float3 o = xmat4_point(cam.c2w, f3(0.f, 0.f, 0.f));   // the camera position: the same for every sample
    unsigned nv = 0, nt = 0, nExt = 0, nShadow = 0;
    const int nc = pixel_candidates(bvh, sp, cam, px, py, o, stack, nv);
    if (nc == 0) {
        const unsigned m = __activemask();
        const unsigned culled = __reduce_add_sync(m, s1 - s0);
        if ((threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(&ctr->rays_culled, (unsigned long long)culled);
        if (!sp.has_env()) return;   // black environment: nothing to add either
    } else if (nc > 0) {
        const unsigned m = __activemask();
        const unsigned n = __reduce_add_sync(m, s1 - s0);
        if ((threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(&ctr->rays_candidates, (unsigned long long)n);
    }
    // the candidates move out of the stack array: the traversals primary_sample may run (light shadow rays, specular
    // continuation) use it from the bottom with no depth limit of their own, and an LBVH can be up to 95 levels deep
    int cand[kMaxCandidates], bound[kMaxCandidates];
    for (int k = 0; k < nc; ++k) {
        cand[k] = stack[kStackSize - 1 - k];
        bound[k] = stack[kStackSize - 1 - kMaxCandidates - k];
    }
    for (uint32_t sl = s0; sl < s1; ++sl)
        primary_sample<COUNT, NEE, MODE>(bvh, sp, cam, q, blockIdx.x & (kSegments - 1), px, py, pixel, sampleBase + sl * sampleStride, o, nc, cand, bound, 1,
                                   stack, aR, aG, aB, nt, nExt, nShadow, ctr);
    if (COUNT) flush_trav_counters(nv, nt, ctr);
    flush_ray_counters(nExt, nShadow, ctr);
    unsigned long long* dst = acc + 3 * (size_t)pixel;
    if (aR) atomicAdd(dst + 0, (unsigned long long)aR);
    if (aG) atomicAdd(dst + 1, (unsigned long long)aG);
    if (aB) atomicAdd(dst + 2, (unsigned long long)aB);
}

// Form B (32 or more samples per pixel in this launch): warp = (8x4 pixel tile, group of `spt` local samples).
// Each lane first walks the tree with ITS pixel's pyramid and leaves the candidates in shared memory; then the warp
// takes the tile's pixels that have work a few at a time, kCoopLanes lanes tracing samples of the same pixel
// together.  Lanes then share candidates, materials and (mostly) control flow, a pixel that needs no work costs
// the warp nothing, and no lane waits long for a neighbour whose pixel happens to be the expensive one.
constexpr int kCoopCandidates = kMaxCandidates < 16 ? kMaxCandidates : 16;
#ifndef TPT_COOP_LANES
#define TPT_COOP_LANES 32
#endif
// Lanes that share one pixel (power of two, 4..32).  Measured (ms per frame, 8 / 16 / 32 lanes): ball 34.8 / 33.2 / 32.8, box 10.67 /
// 10.54 / 10.42, tir 6.21 / 6.23 / 6.21, light 14.74 / 14.86 / 15.08: since k_primary also traces shadow and mirror rays in
// registers, rays that start at the same pixel's surface points walk the tree together best when the whole warp has one pixel.
constexpr int kCoopLanes = TPT_COOP_LANES;
__device__ __forceinline__ long long group_sum_ll(long long v) {
#pragma unroll
    for (int o = kCoopLanes >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#ifndef TPT_COOP_MIN_BLOCKS
#define TPT_COOP_MIN_BLOCKS 7
#endif
template <bool COUNT, bool NEE, int MODE>
__global__ void __launch_bounds__(128, TPT_COOP_MIN_BLOCKS) k_primary_coop(BvhView bvh, ShadeParams sp, CameraParams cam, int W, int H, uint32_t nPadded,
                                                      uint32_t kSamples, uint32_t spt, uint32_t sampleBase, uint32_t sampleStride, QueueOut q,
                                                      unsigned long long* __restrict__ acc, DeviceCounters* ctr) {
    __shared__ int sCand[4][kCoopCandidates][32];    // [warp][k][lane]: leaf k of the lane's pixel
    __shared__ int sBound[4][kCoopCandidates][32];   //                   its lower bound of t (float bits)
    __shared__ TileWalk sTile[4];                    // [warp]: frontier and leaf list of the tile walk
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPadded * ((kSamples + spt - 1) / spt)) return;    // whole warps: nPadded is a multiple of 32
    if (i == 0) atomicAdd(&ctr->rays_primary, (unsigned long long)W * (unsigned long long)H * kSamples);
    const uint32_t g = i / nPadded, p = i - g * nPadded;        // g is the same for the whole warp
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    int px, py;
    const bool inside = tile_decode(p, W, H, px, py);
    const uint32_t s0 = g * spt, s1 = min(kSamples, s0 + spt);
    int stack[kStackSize];
    stack[0] = kStackSentinel;
    const float3 o = xmat4_point(cam.c2w, f3(0.f, 0.f, 0.f));
    unsigned nv = 0, nt = 0, nExt = 0, nShadow = 0;
    int nc = 0;
    // one cooperative walk with the tile's pyramid, then each pixel filters the tile's leaves (tile_collect); per-pixel walks
    // only when the tile meets too many leaves
    int nTile = -1;
    if (sp.pixelCandidates && !sp.noTileWalk) {
        int tx, ty;
        tile_decode(p & ~31u, W, H, tx, ty);   // lane 0's pixel = the tile's corner
        nTile = tile_collect(bvh, tile_frustum(cam, tx, ty, o), sTile[wib], nv);
    }
    if (inside && nTile != 0) {
        nc = nTile > 0 ? candidates_from_tile(bvh, pixel_frustum(cam, px, py, o), sTile[wib], nTile, stack)
                       : pixel_candidates(bvh, sp, cam, px, py, o, stack, nv);
        if (nc > kCoopCandidates) nc = -1;
        for (int k = 0; k < nc; ++k) {
            sCand[wib][k][lane] = stack[kStackSize - 1 - k];
            sBound[wib][k][lane] = stack[kStackSize - 1 - kMaxCandidates - k];
        }
    }
    __syncwarp();
    {   // statistics: camera samples answered without / with candidate tests
        const unsigned nCulled = __popc(__ballot_sync(0xffffffffu, inside && nc == 0)), nCand = __popc(__ballot_sync(0xffffffffu, nc > 0));
        if (lane == 0 && nCulled) atomicAdd(&ctr->rays_culled, (unsigned long long)nCulled * (s1 - s0));
        if (lane == 0 && nCand) atomicAdd(&ctr->rays_candidates, (unsigned long long)nCand * (s1 - s0));
    }
    // kCoopLanes lanes work on one pixel, 32 / kCoopLanes pixels at a time (the next ones that have work)
    const uint32_t grp = lane / kCoopLanes, sub = lane % kCoopLanes;
    unsigned todo = __ballot_sync(0xffffffffu, inside && (nc != 0 || sp.has_env()));
    // Pixels no sample of which can hit anything only look the environment up.  Where they are many (a tile of sky: three quarters
    // of ball.gltf's tiles) each lane keeps ITS pixel and runs the samples one after the other: the same number of warp iterations
    // as one pixel at a time with 32 lanes, without the per-pixel hand-over (shuffles, the 64-bit group sums, the votes), which
    // at 32 samples per pixel and launch — one iteration per pixel — was half of what such a pixel cost.
    const unsigned culledMask = __ballot_sync(0xffffffffu, inside && nc == 0);
    if (sp.has_env() && 2 * __popc(culledMask) * ((s1 - s0 + 31) / 32) >= (s1 - s0)) {
        if (inside && nc == 0) {
            long long aR = 0, aG = 0, aB = 0;
            const uint32_t pixel = (uint32_t)(px + py * W);
            for (uint32_t sl = s0; sl < s1; ++sl) {   // primary_sample for a ray that misses: jitter, direction, environment (weight 1)
                float ju, jv;
                rng_pair(rng_path_key(sp.seedKey, pixel, sampleBase + sl * sampleStride), 0u, ju, jv);
                const float3 term = env_lookup(sp.env, sp.envf, sp.envW, sp.envH, primary_dir(cam, px, py, ju, jv));
                if (term.x != 0.f || term.y != 0.f || term.z != 0.f) {
                    if (finite3(term)) {
                        aR += to_fixed(term.x);
                        aG += to_fixed(term.y);
                        aB += to_fixed(term.z);
                    } else {
                        atomicAdd(&ctr->dropped_nonfinite, 1ull);
                    }
                }
            }
            unsigned long long* dst = acc + 3 * (size_t)pixel;
            if (aR) atomicAdd(dst + 0, (unsigned long long)aR);
            if (aG) atomicAdd(dst + 1, (unsigned long long)aG);
            if (aB) atomicAdd(dst + 2, (unsigned long long)aB);
        }
        __syncwarp();
        todo &= ~culledMask;
    }
    while (todo) {
        const unsigned src = __fns(todo, 0, (int)grp + 1);   // lane of my group's pixel; 0xffffffff when fewer pixels are left
        const bool have = src != 0xffffffffu;
        const int pl = have ? (int)src : 0;
#pragma unroll
        for (int k = 0; k < 32 / kCoopLanes; ++k) todo &= todo - 1;
        const int ncp = __shfl_sync(0xffffffffu, nc, pl), pxp = __shfl_sync(0xffffffffu, px, pl), pyp = __shfl_sync(0xffffffffu, py, pl);
        const uint32_t pixel = (uint32_t)(pxp + pyp * W);
        long long aR = 0, aG = 0, aB = 0;
        if (have)
            for (uint32_t sl = s0 + sub; sl < s1; sl += kCoopLanes)
                primary_sample<COUNT, NEE, MODE>(bvh, sp, cam, q, blockIdx.x & (kSegments - 1), pxp, pyp, pixel, sampleBase + sl * sampleStride, o, ncp,
                                           &sCand[wib][0][pl], &sBound[wib][0][pl], 32, stack, aR, aG, aB, nt, nExt, nShadow, ctr);
        __syncwarp();
        if (__ballot_sync(0xffffffffu, (aR | aG | aB) != 0)) {
            aR = group_sum_ll(aR);
            aG = group_sum_ll(aG);
            aB = group_sum_ll(aB);
            unsigned long long* dst = acc + 3 * (size_t)pixel;
            if (sub == 0 && aR) atomicAdd(dst + 0, (unsigned long long)aR);
            if (sub == 1 % kCoopLanes && aG) atomicAdd(dst + 1, (unsigned long long)aG);
            if (sub == 2 % kCoopLanes && aB) atomicAdd(dst + 2, (unsigned long long)aB);
        }
    }
    if (COUNT) flush_trav_counters(nv, nt, ctr);
    flush_ray_counters(nExt, nShadow, ctr);
}

// Pre-pass for the batch size: how many pixels of each 8x4 tile can emit a path at all (the same pyramid walk k_primary does,
// so "0 candidates" here is "0 candidates" there).  The queue pools are sized for what the live pixels can emit, not for
// every pixel of the frame: ball.gltf has 22 % live pixels, tir.gltf 10 %, and their 256 spp then fit ONE batch (one
// k_primary + max_depth k_wave launches per frame instead of four of each).
__global__ void __launch_bounds__(128) k_tile_live(BvhView bvh, ShadeParams sp, CameraParams cam, int W, int H, uint32_t nPadded,
                                                   uint8_t* __restrict__ tileLive) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPadded) return;   // whole warps: nPadded is a multiple of 32
    int px, py;
    const bool inside = tile_decode(i, W, H, px, py);
    int stack[kStackSize];
    stack[0] = kStackSentinel;
    const float3 o = xmat4_point(cam.c2w, f3(0.f, 0.f, 0.f));
    // "some leaf box meets the pyramid" is exactly "pixel_candidates(...) != 0" (same plane-box test), and stops at the
    // first such leaf instead of collecting and sorting them all.  Like k_primary_coop: one cooperative walk per tile first.
    __shared__ TileWalk sTile[4];
    TileWalk& tw = sTile[threadIdx.x >> 5];
    int nTile = -1;
    if (!sp.noTileWalk) {
        int tx, ty;
        unsigned nvUnused = 0;
        tile_decode(i & ~31u, W, H, tx, ty);
        nTile = tile_collect(bvh, tile_frustum(cam, tx, ty, o), tw, nvUnused);
    }
    bool hit = false;
    if (inside && nTile < 0) {
        hit = frustum_may_hit(bvh, pixel_frustum(cam, px, py, o), stack);
    } else if (inside && nTile > 0) {
        const PixelFrustum f = pixel_frustum(cam, px, py, o);
        for (int j = 0; j < nTile && !hit; ++j) {
            const int code = tw.leafNode[j];
            if (code < 0) {
                hit = true;
                break;
            }
            const float4* n = bvh.nodes + 4 * (size_t)(code >> 1);
            const float4 xy = __ldg(n + (code & 1)), z = __ldg(n + 2);
            hit = frustum_box(f, xy.x, xy.y, xy.z, xy.w, (code & 1) ? z.z : z.x, (code & 1) ? z.w : z.y);
        }
    }
    const unsigned live = __ballot_sync(0xffffffffu, hit);
    if ((threadIdx.x & 31) == 0) tileLive[i >> 5] = (uint8_t)__popc(live);
}

// ---- wave w >= 1: extension rays towards vertex w + secondary rays of vertex w-1 --------------------------
// Persistent warps.  Incoherent rays finish after very different numbers of steps, and a lane that waits
// for the slowest ray of its warp is a wasted lane, so each lane runs the traversal state machine of
// traverse.cuh and the warp schedules itself:
//   * every iteration the warp votes and executes ONE step kind — the box step for the lanes that sit at an
//     internal node, or the (exactly rounded, expensive) triangle step for the lanes that sit at a leaf —
//     whichever has more lanes ready; the others keep their state and wait;
//   * when kRefillThreshold lanes are idle, finished lanes retire their ray and idle lanes take new items
//     (ray replacement); items come from the global cursor 32-256 at a time and are dealt out warp-locally;
//   * retiring an extension ray only stages (item, hit) in shared memory; once 32 are staged the whole
//     warp shades them together (shade_vertex + emit), so the divergent BSDF code also runs on full warps.
// How many lanes must be idle before the warp stops stepping to retire and refill.  A refill phase costs every lane of the warp
// (votes, staging, item loads), so it pays to wait until it serves many lanes — as long as the idle lanes are not too many for
// too long.  Measured (8 / 16 / 20 / 24 idle lanes; ms per frame): ball 31.5 / 30.3 / 29.8 / 29.9, tir 6.26 / 5.77 / 5.68 / 5.53,
// box 10.34 / 10.18 / 10.27 / 10.71; 10 M-triangle soup, Mrays/s primary | secondary: 2066 | 1312, 2250 | 1299, 2361 | 1263,
// 2401 | 1225 — coherent rays want a late refill, incoherent ones an early one.  k_wave: 20; k_trace_rays_persistent, which
// cannot know what it is given: 16.
#ifndef TPT_REFILL_THRESHOLD
#define TPT_REFILL_THRESHOLD 20
#endif
#ifndef TPT_TRACE_REFILL_THRESHOLD
#define TPT_TRACE_REFILL_THRESHOLD 16
#endif
constexpr int kRefillThreshold = TPT_REFILL_THRESHOLD;
constexpr int kTraceRefillThreshold = TPT_TRACE_REFILL_THRESHOLD;
#ifndef TPT_KEEP_NUM
#define TPT_KEEP_NUM 2
#endif
// a step kind is repeated without a re-vote while kKeepNum / 4 of its lanes stay in that state (2 / 3 / 4 of 4, ms per frame:
// box 10.10 / 10.29 / 11.13, ball 29.67 / 30.01 / 30.54; soup within 1 %)
constexpr int kKeepNum = TPT_KEEP_NUM;
constexpr uint32_t kGrabMax = 256;
constexpr int kStage = 64;
constexpr uint32_t kNoItem = 0xffffffffu;

template <bool NEE>
__device__ __forceinline__ void shade_staged(const BvhView& bvh, const ShadeParams& sp, const float4* __restrict__ extIn, const QueueOut& q,
                                             unsigned long long* __restrict__ acc, DeviceCounters* ctr, const uint32_t* myIdx,
                                             const float4* myHit, uint32_t count, int depth) {
    const uint32_t lane = threadIdx.x & 31;
    if (lane < count) {
        const PathRec p = load_path(extIn, myIdx[lane]);
        const float4 hv = myHit[lane];
        HitRec h;
        h.t = hv.x;
        h.u = hv.y;
        h.v = hv.z;
        h.leaf = __float_as_int(hv.w);
        h.prim = -1;
        float3 term;
        Vertex vx;
        const uint32_t ahead = NEE ? 0u : p.w;
        if (shade_vertex<NEE>(bvh, sp, depth + (int)ahead, p.o, p.d, p.T, NEE ? p.w : 0u, p.pixel, p.sample, h, term, vx, ctr))
            emit_vertex<NEE, true>(sp, q, 0u, vx, p.pixel, p.sample, ahead, ctr);
        else accum_add(acc, p.pixel, term, ctr);
    }
    __syncwarp();
}

#ifndef TPT_WAVE_MIN_BLOCKS
#define TPT_WAVE_MIN_BLOCKS 8
#endif
// SEGIN: the input level is the segmented one k_primary wrote (wave 1); k_wave itself appends to segment 0 only — its
// warps issue one atomic per 32 shaded vertices, far below what one counter sustains — so the levels it produces
// are plain queues that use the whole pool.
// One level (= wave `depth`) of the persistent scheduler; every thread of the block calls it.
// LOOP: called level after level inside one kernel (k_wave_loop) — the queues it reads were written earlier in the SAME launch,
// so their loads must not use the non-coherent path (ld.global.nc is only valid for data that is read-only during the launch).
template <bool COUNT, bool NEE, bool SEGIN, bool LOOP = false>
__device__ __forceinline__ void wave_level(const BvhView& bvh, const ShadeParams& sp, int depth, const float4* __restrict__ extIn,
                                           const float4* __restrict__ secIn, void* __restrict__ levelIn, const QueueOut& q,
                                           unsigned long long* __restrict__ acc, DeviceCounters* ctr) {
    __shared__ uint32_t sIdx[4][kStage];
    __shared__ float4 sHit[4][kStage];
    // The level's items form one virtual queue: the extension rays of segment 0, 1, ... then the secondary rays of
    // segment 0, 1, ...; sExtPre / sSecPre hold the exclusive prefix sums of the segment counts.
    __shared__ uint32_t sExtPre[SEGIN ? kSegments + 1 : 1], sSecPre[SEGIN ? kSegments + 1 : 1], sDirect;
    if (!SEGIN) {
        if (threadIdx.x == 0) {
            const unsigned long long w = seg_level(levelIn, 0)->counts;
            sExtPre[0] = min((uint32_t)w, q.extCap);
            sSecPre[0] = min((uint32_t)(w >> 32), q.secCap);
            sDirect = seg_level(levelIn, 0)->n_direct;
        }
    } else if (threadIdx.x < 32) {
        static_assert(kSegments % 32 == 0 || kSegments < 32, "kSegments: a power of two");
        constexpr int per = kSegments >= 32 ? kSegments / 32 : 1;
        const uint32_t l = threadIdx.x;
        uint32_t e[per], c[per], sumE = 0, sumS = 0, sumD = 0;
#pragma unroll
        for (int k = 0; k < per; ++k) {
            const uint32_t seg = l * per + k;
            unsigned long long w = 0;
            uint32_t d = 0;
            if (seg < (uint32_t)kSegments) {
                w = seg_level(levelIn, seg)->counts;
                d = seg_level(levelIn, seg)->n_direct;
            }
            e[k] = min((uint32_t)w, q.extSegCap);
            c[k] = min((uint32_t)(w >> 32), q.secSegCap);
            sumE += e[k];
            sumS += c[k];
            sumD += d;
        }
        uint32_t incE = sumE, incS = sumS;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, incE, o), b = __shfl_up_sync(0xffffffffu, incS, o);
            if (l >= (uint32_t)o) {
                incE += a;
                incS += b;
            }
        }
        uint32_t runE = incE - sumE, runS = incS - sumS;
#pragma unroll
        for (int k = 0; k < per; ++k) {
            const uint32_t seg = l * per + k;
            if (seg < (uint32_t)kSegments) {
                sExtPre[seg] = runE;
                sSecPre[seg] = runS;
            }
            runE += e[k];
            runS += c[k];
        }
        sumD = __reduce_add_sync(0xffffffffu, sumD);
        if (l == 31) {
            sExtPre[kSegments] = incE;
            sSecPre[kSegments] = incS;
            sDirect = sumD;
        }
    }
    __syncthreads();
    const uint32_t nExt = sExtPre[SEGIN ? kSegments : 0], nSec = sSecPre[SEGIN ? kSegments : 0], nDirect = sDirect;
    const uint32_t total = nExt + nSec;
    if (total == 0u) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&ctr->rays_extension, (unsigned long long)nExt);
        atomicAdd(&ctr->rays_direct, (unsigned long long)nDirect);
        atomicAdd(&ctr->rays_shadow, (unsigned long long)(nSec - nDirect));
        atomicAdd(&ctr->rays_queued, (unsigned long long)total);
    }
    uint32_t* cursor = &seg_level(levelIn, 0)->cursor;
    const uint32_t lane = threadIdx.x & 31, ltMask = (1u << lane) - 1u;
    // items taken from the global cursor at a time: large for big waves (few atomics), 32 for small ones so
    // that every resident warp gets work
    const uint32_t kGrab = min(max((total / (gridDim.x * 8u)) & ~31u, 32u), kGrabMax);
    uint32_t* myIdx = sIdx[threadIdx.x >> 5];
    float4* myHit = sHit[threadIdx.x >> 5];
    // local-memory stack here: the shared-memory one (HybridStack, used by k_trace_rays_persistent) measured 2-4 % SLOWER on
    // the bundled scenes' waves (its depth test costs issue slots, and these waves are as much issue- as data-pipe-bound)
    int stackArr[kStackSize];
    stackArr[0] = kStackSentinel;
    const LocalStack stack{stackArr};
    TravState ts;
    ts.cur = kStackSentinel;
    ts.sp = 1;
#if TPT_STACK_TOP_SM
    ts.top = kNoTop;
#endif
    uint32_t item = kNoItem, kind = 0;   // kind: 0 extension, 1 shadow (any hit), 2 direct
    uint32_t lo = 0, hi = 0;             // warp-local range of items not yet dealt out
    uint32_t gBase = 0, gEnd = 0;        // SEGIN: virtual index -> pool slot for the run of one segment the current grab starts in
    bool more = true;                    // the global cursor may still have items
    uint32_t nStaged = 0;
    unsigned nv = 0, nt = 0;

    while (true) {
        const bool atNode = ts.cur >= 0;
        const bool atLeaf = ts.cur < 0 && ts.cur != kStackSentinel;
        const unsigned mNode = __ballot_sync(0xffffffffu, atNode), mLeaf = __ballot_sync(0xffffffffu, atLeaf);
        const unsigned mBusy = mNode | mLeaf;
        if (mBusy == 0u || (more && 32 - __popc(mBusy) >= kRefillThreshold)) {
            // ---- retire finished rays
            const bool fin = item != kNoItem && !atNode && !atLeaf;
            const bool finExt = fin && kind == 0u;
            const unsigned mExt = __ballot_sync(0xffffffffu, finExt);
            if (finExt) {
                const uint32_t slot = nStaged + __popc(mExt & ltMask);
                myIdx[slot] = item;
                myHit[slot] = make_float4(ts.best.t, ts.best.u, ts.best.v, __int_as_float(ts.best.leaf));
            }
            nStaged += __popc(mExt);
            if (fin && kind != 0u) {
                float e = 1.0f;
                bool add = ts.best.leaf < 0;   // shadow ray: unoccluded
                if (kind == 2u) {              // direct ray: emission of what it hit
                    add = false;
                    if (ts.best.leaf >= 0) {
                        const int mtlIdx = __float_as_int(__ldg(bvh.tri_geom + 3 * (size_t)ts.best.leaf + 2).z);
                        e = __ldg(sp.mtl + 2 * mtlIdx).w;
                        add = e != 0.0f;
                    }
                }
                if (add) {
                    const size_t i = (size_t)item;
                    const float4 b = __ldcs(secIn + 3 * i + 1), c = __ldcs(secIn + 3 * i + 2);
                    accum_add(acc, __float_as_uint(c.y), f3(b.z, b.w, c.x) * e, ctr);
                }
            }
            if (fin) item = kNoItem;
            __syncwarp();
            if (nStaged >= 32u) {
                shade_staged<NEE>(bvh, sp, extIn, q, acc, ctr, myIdx, myHit, 32u, depth);
                const uint32_t r = nStaged - 32u;
                uint32_t ti = 0;
                float4 th = make_float4(0.f, 0.f, 0.f, 0.f);
                if (lane < r) {
                    ti = myIdx[32 + lane];
                    th = myHit[32 + lane];
                }
                __syncwarp();
                if (lane < r) {
                    myIdx[lane] = ti;
                    myHit[lane] = th;
                }
                __syncwarp();
                nStaged = r;
            }
            // ---- deal new items to the idle lanes
            if (more && lo == hi) {
                if (lane == 0) lo = atomicAdd(cursor, kGrab);
                lo = __shfl_sync(0xffffffffu, lo, 0);
                hi = min(lo + kGrab, total);
                if (lo >= total) {
                    more = false;
                    lo = hi = 0;
                } else if (SEGIN) {
                    const bool isExt = lo < nExt;
                    const uint32_t v = isExt ? lo : lo - nExt;
                    const uint32_t* pre = isExt ? sExtPre : sSecPre;
                    uint32_t seg = 0;
#pragma unroll
                    for (int step = kSegments >> 1; step > 0; step >>= 1)
                        if (pre[seg + step] <= v) seg += step;
                    gBase = seg * (isExt ? q.extSegCap : q.secSegCap) - pre[seg] - (isExt ? 0u : nExt);   // item = vi + gBase (mod 2^32)
                    gEnd = pre[seg + 1] + (isExt ? 0u : nExt);
                }
            }
            if (more) {
                const bool empty = item == kNoItem;
                const unsigned mEmpty = __ballot_sync(0xffffffffu, empty);
                const uint32_t avail = hi - lo, rank = __popc(mEmpty & ltMask);
                if (empty && rank < avail) {
                    // virtual index -> (segment, offset) -> slot in the extension or secondary pool
                    const uint32_t vi = lo + rank;
                    const bool isExt = vi < nExt;
                    const uint32_t v = isExt ? vi : vi - nExt;
                    item = v;
                    if (SEGIN) {
                        if (vi < gEnd) {   // same run of one segment as the first item of this grab (searched once per grab)
                            item = vi + gBase;
                        } else {
                            const uint32_t* pre = isExt ? sExtPre : sSecPre;
                            uint32_t seg = 0;
#pragma unroll
                            for (int step = kSegments >> 1; step > 0; step >>= 1)
                                if (pre[seg + step] <= v) seg += step;
                            item = seg * (isExt ? q.extSegCap : q.secSegCap) + (v - pre[seg]);
                        }
                    }
                    const float4* src = (isExt ? extIn : secIn) + 3 * (size_t)item;
#ifdef TPT_REFILL_LDCS
                    const float4 a = __ldcs(src), b = __ldcs(src + 1);
                    kind = isExt ? 0u : 1u + __float_as_uint(__ldcs(reinterpret_cast<const float*>(src + 2) + 3));
#else
                    // cached loads: the rest of the item is read again when the ray retires (shade_staged / accumulate)
                    const float4 a = LOOP ? __ldcg(src) : __ldg(src), b = LOOP ? __ldcg(src + 1) : __ldg(src + 1);
                    const float* kp = reinterpret_cast<const float*>(src + 2) + 3;
                    kind = isExt ? 0u : 1u + __float_as_uint(LOOP ? __ldcg(kp) : __ldg(kp));
#endif
                    trav_begin(ts, bvh, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y));
                }
                lo += min((uint32_t)__popc(mEmpty), avail);
            }
            if (!more && __ballot_sync(0xffffffffu, item != kNoItem) == 0u) break;
            continue;
        }
        if (__popc(mNode) >= __popc(mLeaf)) {
            // box steps for the lanes at internal nodes; keep going without a full re-vote while at least
            // three quarters of them are still descending (coherent rays stay in this loop for many steps)
            const int keep = max(1, (kKeepNum * __popc(mNode)) >> 2);
            do {
                if (ts.cur >= 0) {
                    trav_node_step(ts, bvh, stack);
                    if (COUNT) ++nv;
                }
            } while (__popc(__ballot_sync(0xffffffffu, ts.cur >= 0)) >= keep);
        } else {
            const int keep = max(1, (kKeepNum * __popc(mLeaf)) >> 2);
            do {
                if (ts.cur < 0 && ts.cur != kStackSentinel) {
                    trav_leaf_step(ts, bvh, kind == 1u, stack);
                    if (COUNT) ++nt;
                }
            } while (__popc(__ballot_sync(0xffffffffu, ts.cur < 0 && ts.cur != kStackSentinel)) >= keep);
        }
    }
    if (nStaged) shade_staged<NEE>(bvh, sp, extIn, q, acc, ctr, myIdx, myHit, nStaged, depth);
    if (COUNT) {
        const unsigned sv = __reduce_add_sync(0xffffffffu, nv), st = __reduce_add_sync(0xffffffffu, nt);
        if (lane == 0) {
            atomicAdd(&ctr->node_visits, (unsigned long long)sv);
            atomicAdd(&ctr->tri_tests, (unsigned long long)st);
        }
    }
}

template <bool COUNT, bool NEE, bool SEGIN>
__global__ void __launch_bounds__(128, TPT_WAVE_MIN_BLOCKS) k_wave(BvhView bvh, ShadeParams sp, const float4* __restrict__ extIn, const float4* __restrict__ secIn,
                                              void* __restrict__ levelIn, QueueOut q, unsigned long long* __restrict__ acc,
                                              DeviceCounters* ctr) {
    wave_level<COUNT, NEE, SEGIN>(bvh, sp, sp.depth, extIn, secIn, levelIn, q, acc, ctr);
}

// A small level finished path by path.  When a level holds fewer items than the grid has threads, a level-synchronous wave costs
// what its SLOWEST ray costs — measured on ball.gltf: ~80 us for each of waves 2..8 whatever the item count (the survivors are
// the grazing rays, the longest walks there are), 0.54 ms per batch that does not shrink with the samples per GPU.  Here each
// thread takes one item of the level and follows it to the end of its path in registers: closest hit, shade, the vertex's
// shadow / "direct" rays traced at once, next bounce.  The time becomes that of the longest single PATH, not the sum over levels
// of the longest ray.  Every term added is the one the queued form adds (same arithmetic, same RNG dimensions, exact integer
// accumulation), so the accumulator bits are unchanged.
template <bool COUNT, bool NEE>
__device__ __forceinline__ void path_tail(const BvhView& bvh, const ShadeParams& sp, int w, const float4* extIn, const float4* secIn, uint32_t nExt,
                                          uint32_t nSec, uint32_t nDirect, unsigned long long* __restrict__ acc, DeviceCounters* ctr) {
    const uint32_t total = nExt + nSec;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&ctr->rays_extension, (unsigned long long)nExt);
        atomicAdd(&ctr->rays_direct, (unsigned long long)nDirect);
        atomicAdd(&ctr->rays_shadow, (unsigned long long)(nSec - nDirect));
        atomicAdd(&ctr->rays_queued, (unsigned long long)total);
    }
    int stack[kStackSize];
    stack[0] = kStackSentinel;
    unsigned nExtra = 0, nShadowExtra = 0, nDirectExtra = 0;   // rays traced here that never saw a queue
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        if (i >= nExt) {   // a secondary ray of vertex w-1
            const size_t j = (size_t)(i - nExt);
            const float4 a = __ldcg(secIn + 3 * j), b = __ldcg(secIn + 3 * j + 1), c = __ldcg(secIn + 3 * j + 2);
            const uint32_t kind = __float_as_uint(c.w);
            const HitRec h = traverse<COUNT>(bvh, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), kind == kKindShadow, stack, &ctr->node_visits, &ctr->tri_tests);
            float e = 1.0f;
            bool add = h.leaf < 0;
            if (kind == kKindDirect) {
                add = false;
                if (h.leaf >= 0) {
                    e = __ldg(sp.mtl + 2 * __float_as_int(__ldg(bvh.tri_geom + 3 * (size_t)h.leaf + 2).z)).w;
                    add = e != 0.0f;
                }
            }
            if (add) accum_add(acc, __float_as_uint(c.y), f3(b.z, b.w, c.x) * e, ctr);
            continue;
        }
        const float4 a = __ldcg(extIn + 3 * (size_t)i), b = __ldcg(extIn + 3 * (size_t)i + 1), c = __ldcg(extIn + 3 * (size_t)i + 2);
        float3 o = f3(a.x, a.y, a.z), d = f3(a.w, b.x, b.y), T = f3(b.z, b.w, c.x);
        const uint32_t pixel = __float_as_uint(c.y), sample = __float_as_uint(c.z);
        uint32_t pbBits = NEE ? __float_as_uint(c.w) : 0u;
        int depth = w + (NEE ? 0 : (int)__float_as_uint(c.w));
        while (true) {
            const HitRec h = traverse<COUNT>(bvh, o, d, false, stack, &ctr->node_visits, &ctr->tri_tests);
            float3 term;
            Vertex vx;
            if (!shade_vertex<NEE>(bvh, sp, depth, o, d, T, pbBits, pixel, sample, h, term, vx, ctr)) {
                accum_add(acc, pixel, term, ctr);
                break;
            }
            for (int l = 0; l < sp.nLights; ++l) {   // what emit_vertex queues and the next wave adds, in the same arithmetic
                float3 ldir, lrad;
                light_incoming(sp.lights[l], vx.P, ldir, lrad);
                const float3 contrib = vx.Tn * (vx.base * lrad);
                ++nShadowExtra;
                if (traverse<COUNT>(bvh, vx.P, ldir, true, stack, &ctr->node_visits, &ctr->tri_tests).leaf < 0) accum_add(acc, pixel, contrib * 1.0f, ctr);
            }
            if (vx.direct) {
                ++nDirectExtra;
                const HitRec hd = traverse<COUNT>(bvh, vx.P, vx.ddir, false, stack, &ctr->node_visits, &ctr->tri_tests);
                if (hd.leaf >= 0) {
                    const float e = __ldg(sp.mtl + 2 * __float_as_int(__ldg(bvh.tri_geom + 3 * (size_t)hd.leaf + 2).z)).w;
                    if (e != 0.0f) accum_add(acc, pixel, vx.Tn * e, ctr);
                }
            }
            if (NEE && vx.nee) {
                ++nShadowExtra;
                if (traverse<COUNT>(bvh, vx.P, vx.ldir, true, stack, &ctr->node_visits, &ctr->tri_tests).leaf < 0) accum_add(acc, pixel, vx.lw * 1.0f, ctr);
            }
            if (!vx.next) break;
            o = vx.P;
            d = vx.ndir;
            T = vx.Tn;
            pbBits = NEE ? __float_as_uint(vx.pb) : 0u;
            ++depth;
            ++nExtra;
        }
    }
    const unsigned m = __activemask();
    const unsigned se = __reduce_add_sync(m, nExtra), ss = __reduce_add_sync(m, nShadowExtra), sd = __reduce_add_sync(m, nDirectExtra);
    if ((threadIdx.x & 31) == __ffs(m) - 1) {
        if (se) atomicAdd(&ctr->rays_extension, (unsigned long long)se);
        if (ss) atomicAdd(&ctr->rays_shadow, (unsigned long long)ss);
        if (sd) atomicAdd(&ctr->rays_direct, (unsigned long long)sd);
    }
}

// Waves wFirst..wLast in ONE cooperative launch: the same scheduler level after level, separated by a grid-wide barrier
// instead of a kernel boundary, and the loop ends as soon as a level is empty.  The deep waves of most frames hold a
// tiny share of the paths (ball.gltf: waves 2-8 together 0.15 % of wave 1) and cost a launch each — seven launches of
// 0.1-0.2 ms per batch, which is what did not shrink when the samples were spread over 8 GPUs.
// Same items, same order-independent integer accumulation: the accumulator bits do not change.
struct WaveLoopParams {
    float4* ext[2];
    float4* sec[2];
    char* levels;               // level w's counter blocks at levels + w * levelBytes
    uint32_t levelBytes;
    int wFirst, wLast;
    uint32_t extSegCap, secSegCap, extCap, secCap;
    uint32_t pathTailMax;       // a level with at most this many items is finished path by path (path_tail)
};
template <bool COUNT, bool NEE>
__global__ void __launch_bounds__(128, 6) k_wave_loop(BvhView bvh, ShadeParams sp, WaveLoopParams lp, unsigned long long* __restrict__ acc,
                                                                      DeviceCounters* ctr) {
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    for (int w = lp.wFirst; w <= lp.wLast; ++w) {
        const int in = w & 1, out = in ^ 1;
        void* levelIn = lp.levels + (size_t)w * lp.levelBytes;
        // written by the previous level (or launch) and final since the barrier: the same value in every block
        const unsigned long long counts = __ldcg(&seg_level(levelIn, 0)->counts);
        if (counts == 0ull) break;
        const uint32_t nExt = min((uint32_t)counts, lp.extCap), nSec = min((uint32_t)(counts >> 32), lp.secCap);
        if (nExt + nSec <= lp.pathTailMax) {   // few items left: finish them path by path, no further level
            path_tail<COUNT, NEE>(bvh, sp, w, lp.ext[in], lp.sec[in], nExt, nSec, __ldcg(&seg_level(levelIn, 0)->n_direct), acc, ctr);
            break;
        }
        const QueueOut q{lp.ext[out], lp.sec[out], lp.levels + (size_t)(w + 1) * lp.levelBytes, lp.extSegCap, lp.secSegCap, lp.extCap, lp.secCap};
        wave_level<COUNT, NEE, false, true>(bvh, sp, w, lp.ext[in], lp.sec[in], levelIn, q, acc, ctr);
        if (w < lp.wLast) grid.sync();
    }
}

// ---- tpt_trace_rays, persistent form: the scheduler of k_wave without the shading ------------------------
template <bool COUNT, bool WIDE = false>
__global__ void __launch_bounds__(128, 10) k_trace_rays_persistent(BvhView bvh, const tpt_ray* __restrict__ rays, unsigned long long n,
                                                                   tpt_hit* __restrict__ hits, int anyHit, unsigned long long* cursor,
                                                                   DeviceCounters* ctr) {
    const uint32_t lane = threadIdx.x & 31, ltMask = (1u << lane) - 1u;
    const unsigned long long perWarp = n / ((unsigned long long)gridDim.x * 8ull);
    const unsigned long long grab = perWarp >= kGrabMax ? kGrabMax : (perWarp >= 32ull ? (perWarp & ~31ull) : 32ull);
#if TPT_SMEM_STACK > 0
    __shared__ int sStack[kSmemStack][128];
    int stackDeep[(WIDE ? kWideStackSize : kStackSize) - kSmemStack];   // wide nodes push up to three children per visit
    const HybridStack stack{&sStack[0][threadIdx.x], stackDeep};
    sStack[0][threadIdx.x] = kStackSentinel;
#else
    int stackArr[WIDE ? kWideStackSize : kStackSize];
    stackArr[0] = kStackSentinel;
    const LocalStack stack{stackArr};
#endif
    TravState ts;
    ts.cur = kStackSentinel;
    ts.sp = 1;
#if TPT_STACK_TOP_SM
    ts.top = kNoTop;
#endif
    unsigned long long item = ~0ull, lo = 0, hi = 0;
    bool more = true;
    unsigned nv = 0, nt = 0;
    while (true) {
        const unsigned mNode = __ballot_sync(0xffffffffu, ts.cur >= 0);
        const unsigned mLeaf = __ballot_sync(0xffffffffu, ts.cur < 0 && ts.cur != kStackSentinel);
        const unsigned mBusy = mNode | mLeaf;
        if (mBusy == 0u || (more && 32 - __popc(mBusy) >= kTraceRefillThreshold)) {
            if (item != ~0ull && !((mBusy >> lane) & 1u)) {   // retire: write the hit
                reinterpret_cast<float4*>(hits)[item] = make_float4(__int_as_float(ts.best.prim), ts.best.t, ts.best.u, ts.best.v);
                item = ~0ull;
            }
            if (more && lo == hi) {
                if (lane == 0) lo = atomicAdd(cursor, grab);
                lo = __shfl_sync(0xffffffffu, lo, 0);
                hi = min(lo + grab, n);
                if (lo >= n) {
                    more = false;
                    lo = hi = 0;
                }
            }
            if (more) {
                const bool empty = item == ~0ull;
                const unsigned mEmpty = __ballot_sync(0xffffffffu, empty);
                const unsigned long long avail = hi - lo;
                const uint32_t rank = __popc(mEmpty & ltMask);
                if (empty && rank < avail) {
                    item = lo + rank;
                    const float2* r = reinterpret_cast<const float2*>(rays + item);
                    const float2 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2);
                    trav_begin(ts, bvh, f3(a.x, a.y, b.x), f3(b.y, c.x, c.y));
                }
                lo += min((unsigned long long)__popc(mEmpty), avail);
            }
            if (!more && __ballot_sync(0xffffffffu, item != ~0ull) == 0u) break;
            continue;
        }
        if (__popc(mNode) >= __popc(mLeaf)) {
            const int keep = max(1, (kKeepNum * __popc(mNode)) >> 2);
            do {
                if (ts.cur >= 0) {
                    if (WIDE) trav_node_step4(ts, bvh.nodes4, stack);
                    else trav_node_step(ts, bvh, stack);
                    if (COUNT) ++nv;
                }
            } while (__popc(__ballot_sync(0xffffffffu, ts.cur >= 0)) >= keep);
        } else {
            const int keep = max(1, (kKeepNum * __popc(mLeaf)) >> 2);
            do {
                if (ts.cur < 0 && ts.cur != kStackSentinel) {
                    trav_leaf_step(ts, bvh, anyHit != 0, stack);
                    if (COUNT) ++nt;
                }
            } while (__popc(__ballot_sync(0xffffffffu, ts.cur < 0 && ts.cur != kStackSentinel)) >= keep);
        }
    }
    if (COUNT) {
        const unsigned sv = __reduce_add_sync(0xffffffffu, nv), st = __reduce_add_sync(0xffffffffu, nt);
        if (lane == 0) {
            atomicAdd(&ctr->node_visits, (unsigned long long)sv);
            atomicAdd(&ctr->tri_tests, (unsigned long long)st);
        }
    }
}

// ---- accumulation buffer utilities ---------------------------------------------------------------
__global__ void k_accum_add(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ src, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] += src[i];
}
__global__ void k_accum_to_f32(const unsigned long long* __restrict__ acc, float* __restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = (float)((double)(long long)acc[i] * (1.0 / 68719476736.0));
}

// copyToFB, src/path_tracer.cu:451-471 (to_uchar / accum_to_radiance: common.cuh): B,G,R order, alpha untouched, row H-1-y.
__global__ void __launch_bounds__(256) k_resolve(const unsigned long long* __restrict__ acc, int W, int H, float invSpp, uchar4* __restrict__ fb) {
    const uint32_t n = (uint32_t)W * (uint32_t)H;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int x = (int)(i % (uint32_t)W), y = (int)(i / (uint32_t)W);
        const unsigned long long* p = acc + 3 * (size_t)i;
        const float r = accum_to_radiance(p[0], invSpp), g = accum_to_radiance(p[1], invSpp), b = accum_to_radiance(p[2], invSpp);
        uchar4* dst = fb + (size_t)(H - y - 1) * W + x;
        uchar4 px = *dst;
        px.x = to_uchar(b);
        px.y = to_uchar(g);
        px.z = to_uchar(r);
        *dst = px;
    }
}

}   // namespace tpt

using namespace tpt;

static inline unsigned grid_for(size_t n, int block, unsigned cap) {
    size_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    return (unsigned)(g < cap ? g : cap);
}

static CameraParams make_camera(const tpt_camera* cam, int W, int H) {
    CameraParams c;
    memcpy(c.c2w, cam->c2w, sizeof c.c2w);
    const float tanHalf = tanf(cam->vfov * 0.5f);
    c.sensorH = 2.0f * tanHalf;
    c.sensorW = cam->aspect * c.sensorH;
    c.invW = 1.0f / (float)W;
    c.invH = 1.0f / (float)H;
    return c;
}

template <typename K>
static int resident_blocks(K kernel, int block) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, block, 0) != cudaSuccess || n < 1) n = 1;
    return n;
}

int tpt_primary_rays_impl(tpt_context* ctx, const tpt_camera* cam, int W, int H, int jitter, uint64_t seed, tpt_ray* d_rays) {
    k_primary_rays<<<grid_for((size_t)W * H, 256, kNumSMs * 8), 256, 0, ctx->stream>>>(make_camera(cam, W, H), W, H, jitter, rng_seed_key(seed), d_rays);
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += 1;
    return TPT_OK;
}

int tpt_trace_rays_impl(tpt_scene* s, const tpt_ray* d_rays, size_t n, tpt_hit* d_hits, int anyHit, int counted) {
    tpt_context* ctx = s->ctx;
    const BvhView v = s->view();
    static const bool simple = getenv("TPT_TRACE_SIMPLE") != nullptr;   // tuning switch: static grid-stride kernel
    if (simple) {
        const unsigned grid = grid_for(n, 128, kNumSMs * 16);
        if (anyHit) {
            if (counted) k_trace_rays<true, true><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
            else k_trace_rays<true, false><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
        } else {
            if (counted) k_trace_rays<false, true><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
            else k_trace_rays<false, false><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
        }
    } else {
        if (!ctx->trace_cursor) TPT_CUDA(cudaMalloc(&ctx->trace_cursor, sizeof(unsigned long long)));
        if (ctx->trace_grid == 0) {
            ctx->trace_grid = kNumSMs * resident_blocks(k_trace_rays_persistent<false, false>, 128);
            ctx->trace_grid_counted = kNumSMs * resident_blocks(k_trace_rays_persistent<true, false>, 128);
            ctx->trace_grid_wide = kNumSMs * resident_blocks(k_trace_rays_persistent<false, true>, 128);
            ctx->trace_grid_wide_counted = kNumSMs * resident_blocks(k_trace_rays_persistent<true, true>, 128);
        }
        TPT_CUDA(cudaMemsetAsync(ctx->trace_cursor, 0, sizeof(unsigned long long), ctx->stream));
        static const bool noWide = getenv("TPT_NO_WIDE") != nullptr;   // tuning / check switch: binary nodes even where wide ones exist
        const bool wide = v.nodes4 != nullptr && !noWide && s->n_faces > 1;
        if (wide && counted)
            k_trace_rays_persistent<true, true><<<ctx->trace_grid_wide_counted, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, anyHit, ctx->trace_cursor, ctx->d_counters);
        else if (wide)
            k_trace_rays_persistent<false, true><<<ctx->trace_grid_wide, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, anyHit, ctx->trace_cursor, ctx->d_counters);
        else if (counted)
            k_trace_rays_persistent<true, false><<<ctx->trace_grid_counted, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, anyHit, ctx->trace_cursor, ctx->d_counters);
        else
            k_trace_rays_persistent<false, false><<<ctx->trace_grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, anyHit, ctx->trace_cursor, ctx->d_counters);
    }
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += 1;
    return TPT_OK;
}

constexpr int kMaxLevels = 72;   // waves 0..max_depth (max_depth <= 64), see tpt_render's argument check

// Queue pools: a batch of P camera samples can emit at most P extension rays and P*(L+1) secondary rays
// per wave (every live path emits one ray of each sort and one shadow ray per light).  `paths` = kSegments x the
// segment capacity, which is what the k_primary blocks of one segment can emit (launch_plan below).
static int pool_alloc_failed(tpt_context* ctx, size_t bytes) {
    cudaGetLastError();
    size_t freeB = 0, totalB = 0;
    cudaMemGetInfo(&freeB, &totalB);
    ctx->pool_budget = 0;   // look the free memory up again next time
    set_error("tpt_render: cannot allocate a " + std::to_string(bytes >> 20) + " MiB wavefront queue (" + std::to_string(freeB >> 20) + " MiB free on device " +
              std::to_string(ctx->device) + "); render with a smaller tpt_render_config.paths_in_flight or free device memory");
    return TPT_ERR_NOMEM;
}

static int ensure_pools(tpt_context* ctx, size_t paths, size_t secPerPath) {
    const size_t secNeed = paths * secPerPath;
    if (!ctx->levels) TPT_CUDA(cudaMalloc(&ctx->levels, (size_t)kMaxLevels * kSegments * kLevelStride));
    if (paths > ctx->ext_capacity) {
        TPT_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 2; ++k) {
            if (ctx->q_ext[k]) cudaFree(ctx->q_ext[k]);
            ctx->q_ext[k] = nullptr;
        }
        ctx->ext_capacity = 0;
        for (int k = 0; k < 2; ++k)
            if (cudaMalloc(&ctx->q_ext[k], paths * 3 * sizeof(float4)) != cudaSuccess) return pool_alloc_failed(ctx, paths * 3 * sizeof(float4));
        ctx->ext_capacity = paths;
    }
    if (secNeed > ctx->sec_capacity) {
        TPT_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 2; ++k) {
            if (ctx->q_sec[k]) cudaFree(ctx->q_sec[k]);
            ctx->q_sec[k] = nullptr;
        }
        ctx->sec_capacity = 0;
        for (int k = 0; k < 2; ++k)
            if (cudaMalloc(&ctx->q_sec[k], secNeed * 3 * sizeof(float4)) != cudaSuccess) return pool_alloc_failed(ctx, secNeed * 3 * sizeof(float4));
        ctx->sec_capacity = secNeed;
    }
    return TPT_OK;
}


int tpt_env_build_sampler(tpt_context* ctx);   // env.cu
int tpt_scene_ensure_normals(tpt_scene* s);     // bvh.cu

int tpt_render_impl(tpt_scene* s, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc) {
    tpt_context* ctx = s->ctx;
    cudaStream_t st = ctx->stream;
    // environment next-event estimation needs a map to sample; without one the flag changes nothing
    const bool nee = (cfg->flags & TPT_RENDER_ENV_NEE) != 0 && (ctx->env != nullptr || ctx->envf != nullptr);
    if (nee) TPT_CHECK(tpt_env_build_sampler(ctx));
    TPT_CHECK(tpt_scene_ensure_normals(s));   // the shade step reads the per-leaf normal records
    const int W = cfg->width, H = cfg->height;
    const uint32_t nPadded = tiles_x(W) * tiles_y(H) * 32u;
    const CameraParams cp = make_camera(cam, W, H);
    ShadeParams sp;
    sp.mtl = s->mtl_table;
    sp.lights = s->lights;
    sp.nLights = (int)s->n_lights;
    sp.env = ctx->env;
    sp.envf = ctx->envf;
    sp.envW = ctx->env_w;
    sp.envH = ctx->env_h;
    sp.maxDepth = cfg->max_depth;
    sp.seedKey = rng_seed_key(cfg->seed);
    sp.depth = 0;
    sp.directRays = s->any_emissive ? 1 : 0;
    sp.pixelCandidates = (cfg->flags & TPT_RENDER_NO_PIXEL_CANDIDATES) ? 0 : 1;
    static const bool noTileWalk = getenv("TPT_NO_TILE_WALK") != nullptr;
    sp.noTileWalk = noTileWalk ? 1 : 0;
    sp.roughMetals = (cfg->flags & TPT_RENDER_ROUGH_METALS) ? 1 : 0;
    static const int specChain = getenv("TPT_SPEC_CHAIN") ? atoi(getenv("TPT_SPEC_CHAIN")) : 4;   // tuning switches
    static const int specVote = getenv("TPT_SPEC_VOTE") ? atoi(getenv("TPT_SPEC_VOTE")) : 3;
    sp.continueSpecular = (cfg->flags & TPT_RENDER_NO_SPECULAR_CONTINUATION) ? 0 : specChain;
    sp.continueVote = specVote;
    sp.es = EnvSampler{ctx->env_prob, ctx->env_alias, ctx->env_mass, ctx->env_w, ctx->env_h};
    const BvhView view = s->view();

    // batch budget: as many LIVE camera samples as a 40 GiB queue budget allows (2 sides x (2+L) rays x 48 B each), at most
    // 2^27 — sized for 180 GB of HBM: fewer, larger batches mean fewer (mostly empty) deep-wave launches per frame
    const size_t secPerPath = s->n_lights + 1 + (nee ? 1 : 0);
    // ... and as the device can spare: at most 60 % of what is free (plus what this context's pools already hold), looked up when the
    // context first renders (and again after an allocation has failed) — several contexts on one GPU, or a GPU shared with other
    // work, get smaller batches instead of a failed cudaMalloc
    if (ctx->pool_budget == 0) {
        size_t freeB = 0, totalB = 0;
        TPT_CUDA(cudaMemGetInfo(&freeB, &totalB));
        const size_t have = (ctx->ext_capacity + ctx->sec_capacity) * 2 * 3 * sizeof(float4);
        ctx->pool_budget = std::max<size_t>(std::min<size_t>((size_t)40 << 30, (freeB + have) / 10 * 6), (size_t)64 << 20);
    }
    size_t target = ctx->pool_budget / (96 * (1 + secPerPath));
    if (target > ((size_t)1 << 27)) target = (size_t)1 << 27;
    if (cfg->paths_in_flight > 0) target = (size_t)cfg->paths_in_flight;
    // Live pixels per 8x4 tile (k_tile_live): only they can put rays into the queues.  One small kernel, one 1-byte-per-tile
    // copy and one stream sync per (geometry, camera, resolution); frames that change none of them reuse the result.
    const uint32_t nTiles = nPadded / 32u;
    static const bool noLiveBound = getenv("TPT_NO_LIVE_BOUND") != nullptr;   // tuning switch: size the pools for every pixel
    {
        LiveKey key;
        memset(&key, 0, sizeof key);
        key.scene = s;
        key.epoch = s->geom_epoch;
        key.cam = *cam;
        key.w = W;
        key.h = H;
        key.candidates = sp.pixelCandidates;
        // the pre-pass costs a stream sync: only worth it when the frame would otherwise need more than one batch
        key.measured = (!noLiveBound && cfg->paths_in_flight <= 0 && (size_t)nPadded * (size_t)cfg->spp > target) ? 1 : 0;
        if (ctx->tile_live_host.size() != nTiles || memcmp(&key, &ctx->live_key, sizeof key) != 0) {
            if (nTiles > ctx->tile_live_capacity) {
                TPT_CUDA(cudaStreamSynchronize(st));
                if (ctx->d_tile_live) cudaFree(ctx->d_tile_live);
                ctx->d_tile_live = nullptr;
                ctx->tile_live_capacity = 0;
                TPT_CUDA(cudaMalloc(&ctx->d_tile_live, nTiles));
                ctx->tile_live_capacity = nTiles;
            }
            ctx->tile_live_host.resize(nTiles);
            if (!key.measured) {
                std::fill(ctx->tile_live_host.begin(), ctx->tile_live_host.end(), (uint8_t)32);
            } else {
                k_tile_live<<<(nPadded + 127) / 128, 128, 0, st>>>(view, sp, cp, W, H, nPadded, ctx->d_tile_live);
                ctx->kernel_launches += 1;
                TPT_CUDA(cudaMemcpyAsync(ctx->tile_live_host.data(), ctx->d_tile_live, nTiles, cudaMemcpyDeviceToHost, st));
                TPT_CUDA(cudaStreamSynchronize(st));
            }
            memcpy(&ctx->live_key, &key, sizeof key);   // padding included: the comparison above is a memcmp
        }
    }
    const uint8_t* tileLive = ctx->tile_live_host.data();
    size_t livePixels = 0;
    for (uint32_t t = 0; t < nTiles; ++t) livePixels += tileLive[t];
    // an explicit paths_in_flight keeps its old meaning (camera samples per batch, live or not): the tests use it to force batching
    uint32_t k = (uint32_t)std::min<size_t>(target / (cfg->paths_in_flight > 0 ? (size_t)nPadded : std::max<size_t>(livePixels, 1)), 1u << 20);
    if (k > 32) k &= ~31u;   // whole warps of samples per pixel for k_primary_coop
    if (k < 1) k = 1;
    if (k > (uint32_t)cfg->spp) k = (uint32_t)cfg->spp;
    // launch shape of k_primary for a batch of kk samples per pixel, and the segment capacity it needs: what the live pixels
    // of the blocks of one segment (blockIdx % kSegments) can emit, one extension ray per live pixel and sample
    struct Plan {
        uint32_t spt, groups;
        unsigned grid;
        bool coop;
        size_t segCap;
    };
    auto launch_plan = [&](uint32_t kk) {
        Plan pl;
        // enough threads to fill the machine several times over, else as many samples per thread as possible
        pl.spt = (uint32_t)(((size_t)nPadded * kk) / ((size_t)kNumSMs * 2048 * 4));
        if (pl.spt < 1) pl.spt = 1;
        static const int sptMax = getenv("TPT_SPT_MAX") ? atoi(getenv("TPT_SPT_MAX")) : 0;   // tuning switch
        if (sptMax > 0 && pl.spt > (uint32_t)sptMax) pl.spt = (uint32_t)sptMax;
        static const bool noCoop = getenv("TPT_NO_COOP") != nullptr;                          // tuning switch
        pl.coop = kk >= 32 && !noCoop;   // a warp per pixel tile, 32 samples of one pixel at a time
        if (pl.coop) pl.spt = (pl.spt + 31u) & ~31u;
        if (pl.spt > kk) pl.spt = kk;
        pl.groups = (kk + pl.spt - 1) / pl.spt;
        pl.grid = (unsigned)(((size_t)nPadded * pl.groups + 127) / 128);
        // warp w of the grid = (sample group w / nTiles, tile w % nTiles), block w / 4, segment block % kSegments
        size_t segSum[kSegments];
        for (int q = 0; q < kSegments; ++q) segSum[q] = 0;
        for (uint32_t g = 0; g < pl.groups; ++g) {
            const size_t ns = std::min(kk, (g + 1) * pl.spt) - g * pl.spt;
            for (uint32_t t = 0; t < nTiles; ++t) segSum[(((size_t)g * nTiles + t) >> 2) & (kSegments - 1)] += tileLive[t] * ns;
        }
        pl.segCap = 32;
        for (int q = 0; q < kSegments; ++q) pl.segCap = std::max(pl.segCap, segSum[q]);
        return pl;
    };
    size_t extSegCap = launch_plan(k).segCap;
    if ((uint32_t)cfg->spp % k) {
        const size_t c = launch_plan((uint32_t)cfg->spp % k).segCap;
        if (c > extSegCap) extSegCap = c;
    }
    const size_t secSegCap = extSegCap * secPerPath;
    TPT_REQUIRE(secSegCap * kSegments < ((size_t)1 << 32), "tpt_render: paths_in_flight x (lights + 2) exceeds 2^32 queue slots");
    TPT_CHECK(ensure_pools(ctx, extSegCap * kSegments, secPerPath));

    const bool counted = (cfg->flags & TPT_RENDER_COUNTERS) != 0;
    const bool timed = (cfg->flags & TPT_RENDER_STAGE_TIMES) != 0;
    const size_t nEvents = 3 + (size_t)cfg->max_depth;
    if (timed) {
        while (ctx->stage_events.size() < nEvents) {
            cudaEvent_t e;
            TPT_CUDA(cudaEventCreate(&e));
            ctx->stage_events.push_back(e);
        }
    }
    size_t ev = 0;
    auto mark = [&]() { if (timed) cudaEventRecord(ctx->stage_events[ev++], st); };
    if (ctx->wave_grid[0] == 0) {
        ctx->wave_grid[0] = kNumSMs * min(resident_blocks(k_wave<false, false, false>, 128), resident_blocks(k_wave<false, false, true>, 128));
        ctx->wave_grid[1] = kNumSMs * min(resident_blocks(k_wave<true, false, false>, 128), resident_blocks(k_wave<true, false, true>, 128));
        ctx->wave_grid[2] = kNumSMs * min(resident_blocks(k_wave<false, true, false>, 128), resident_blocks(k_wave<false, true, true>, 128));
        ctx->wave_grid[3] = kNumSMs * min(resident_blocks(k_wave<true, true, false>, 128), resident_blocks(k_wave<true, true, true>, 128));
    }
    const int variant = (counted ? 1 : 0) + (nee ? 2 : 0);
    const unsigned gridWave = ctx->wave_grid[variant];
    // One launch per wave by default.  The cooperative multi-level kernel (k_wave_loop, TPT_WAVE_LOOP=1) was built to save the deep
    // waves' launches and measured: no gain (ball.gltf, 32 spp per launch: 4.68 vs 4.64 ms per frame) — a small level costs the
    // latency of its slowest ray (~80 us), not a launch — and its 888 resident blocks hold the register file while they wait at the
    // grid barrier, which keeps the NEXT frame's kernels (another stream) off the SMs; the per-wave kernels exit at once when a level
    // is empty and let them in.
    static const bool noWaveLoop = getenv("TPT_WAVE_LOOP") == nullptr || getenv("TPT_NO_WAVE_LOOP") != nullptr;
    const void* kWaveLoop = variant == 0 ? (const void*)k_wave_loop<false, false> : variant == 1 ? (const void*)k_wave_loop<true, false>
                          : variant == 2 ? (const void*)k_wave_loop<false, true> : (const void*)k_wave_loop<true, true>;
    if (ctx->wave_loop_grid[variant] == 0) {
        int coop = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device);
        ctx->wave_loop_grid[variant] = !coop ? 1u   // 1 = not available (a real grid is a multiple of 148)
            : kNumSMs * (unsigned)(variant == 0 ? resident_blocks(k_wave_loop<false, false>, 128) : variant == 1 ? resident_blocks(k_wave_loop<true, false>, 128)
                                 : variant == 2 ? resident_blocks(k_wave_loop<false, true>, 128) : resident_blocks(k_wave_loop<true, true>, 128));
    }
    const bool waveLoop = !noWaveLoop && ctx->wave_loop_grid[variant] > 1u;
    // SPEC kernels (specular continuation) only where they pay: the scene has a mirror / glass material, no env NEE, and the
    // tree is shallow.  Following a ray in k_primary saves a queue round trip and k_wave's scheduling (a fixed cost per ray) but
    // walks the tree in lock-step with the warp's other rays; measured: light.gltf (80 triangles) 18.1 -> 14.7 ms, tir.gltf
    // (6) 6.37 -> 6.21 ms, box.gltf (1 932, long divergent walks inside the box) 10.7 -> 11.1 ms, hence the limit.
    static const size_t specMaxFaces = getenv("TPT_SPEC_MAX_FACES") ? (size_t)atoll(getenv("TPT_SPEC_MAX_FACES")) : 512;   // tuning switch
    const bool spec = s->any_specular && !nee && sp.continueSpecular > 0 && s->n_faces <= specMaxFaces;
    // LHERE kernels: the camera vertex's delta-light shadow rays are traced in k_primary (coherent: one light, neighbouring points)
    static const bool noLightsHere = getenv("TPT_NO_LIGHTS_HERE") != nullptr;   // tuning switch
    const bool lhere = s->n_lights > 0 && !nee && !noLightsHere && !(cfg->flags & TPT_RENDER_NO_SPECULAR_CONTINUATION);
    const int mode = (spec ? 1 : 0) | (lhere ? 2 : 0);
    using PrimaryKernel = void (*)(BvhView, ShadeParams, CameraParams, int, int, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, QueueOut,
                                   unsigned long long*, DeviceCounters*);
    static const PrimaryKernel primaryTable[2][4] = {{k_primary<false, false, 0>, k_primary<false, false, 1>, k_primary<false, false, 2>, k_primary<false, false, 3>},
                                                     {k_primary<true, false, 0>, k_primary<true, false, 1>, k_primary<true, false, 2>, k_primary<true, false, 3>}};
    static const PrimaryKernel coopTable[2][4] = {
        {k_primary_coop<false, false, 0>, k_primary_coop<false, false, 1>, k_primary_coop<false, false, 2>, k_primary_coop<false, false, 3>},
        {k_primary_coop<true, false, 0>, k_primary_coop<true, false, 1>, k_primary_coop<true, false, 2>, k_primary_coop<true, false, 3>}};
    const PrimaryKernel kPrimary = nee ? (counted ? k_primary<true, true, 0> : k_primary<false, true, 0>) : primaryTable[counted ? 1 : 0][mode];
    const PrimaryKernel kPrimaryCoop = nee ? (counted ? k_primary_coop<true, true, 0> : k_primary_coop<false, true, 0>) : coopTable[counted ? 1 : 0][mode];
    const auto kWave = variant == 0 ? k_wave<false, false, false> : variant == 1 ? k_wave<true, false, false> : variant == 2 ? k_wave<false, true, false> : k_wave<true, true, false>;
    const auto kWave1 = variant == 0 ? k_wave<false, false, true> : variant == 1 ? k_wave<true, false, true> : variant == 2 ? k_wave<false, true, true> : k_wave<true, true, true>;

    for (uint32_t done = 0; done < (uint32_t)cfg->spp; done += k) {
        const uint32_t kk = ((uint32_t)cfg->spp - done) < k ? ((uint32_t)cfg->spp - done) : k;
        const Plan pl = launch_plan(kk);
        char* levels = static_cast<char*>(ctx->levels);
        const size_t levelBytes = (size_t)kSegments * kLevelStride;
        TPT_CUDA(cudaMemsetAsync(levels, 0, (size_t)(cfg->max_depth + 2) * levelBytes, st));
        ev = 0;
        mark();
        {
            const QueueOut q{ctx->q_ext[1], ctx->q_sec[1], levels + levelBytes, (uint32_t)extSegCap, (uint32_t)secSegCap, (uint32_t)extSegCap, (uint32_t)secSegCap};
            const uint32_t sampleBase = (uint32_t)cfg->sample_offset + done * (uint32_t)cfg->sample_stride;
            if (pl.coop)
                kPrimaryCoop<<<pl.grid, 128, 0, st>>>(view, sp, cp, W, H, nPadded, kk, pl.spt, sampleBase, (uint32_t)cfg->sample_stride, q, acc->data, ctx->d_counters);
            else
                kPrimary<<<pl.grid, 128, 0, st>>>(view, sp, cp, W, H, nPadded, kk, pl.spt, sampleBase, (uint32_t)cfg->sample_stride, q, acc->data, ctx->d_counters);
        }
        mark();
        // wave w traces extension rays towards vertex w and the secondary rays of vertex w-1; the last wave
        // (w == max_depth) only has secondary rays left.
        const int perLevelLast = waveLoop ? 1 : cfg->max_depth;   // wave 1 reads the segmented level k_primary wrote: its own launch
        for (int w = 1; w <= perLevelLast; ++w) {
            const int in = w & 1, out = in ^ 1;
            sp.depth = w;
            const QueueOut q{ctx->q_ext[out], ctx->q_sec[out], levels + (size_t)(w + 1) * levelBytes, (uint32_t)extSegCap, (uint32_t)secSegCap,
                             (uint32_t)(extSegCap * kSegments), (uint32_t)(secSegCap * kSegments)};
            (w == 1 ? kWave1 : kWave)<<<gridWave, 128, 0, st>>>(view, sp, ctx->q_ext[in], ctx->q_sec[in], levels + (size_t)w * levelBytes, q, acc->data, ctx->d_counters);
            mark();
        }
        uint64_t nLaunches = 1 + (uint64_t)perLevelLast;
        if (waveLoop && cfg->max_depth >= 2) {   // waves 2..max_depth: one cooperative launch, a grid barrier between levels
            WaveLoopParams lp;
            lp.ext[0] = ctx->q_ext[0]; lp.ext[1] = ctx->q_ext[1];
            lp.sec[0] = ctx->q_sec[0]; lp.sec[1] = ctx->q_sec[1];
            lp.levels = levels;
            lp.levelBytes = (uint32_t)levelBytes;
            lp.wFirst = 2;
            lp.wLast = cfg->max_depth;
            lp.extSegCap = (uint32_t)extSegCap; lp.secSegCap = (uint32_t)secSegCap;
            lp.extCap = (uint32_t)(extSegCap * kSegments); lp.secCap = (uint32_t)(secSegCap * kSegments);
            static const long pathTailEnv = getenv("TPT_PATH_TAIL_MAX") ? atol(getenv("TPT_PATH_TAIL_MAX")) : -1;   // tuning switch (0 = off)
            // Off by default: measured on ball.gltf at 32 spp per launch the path-by-path finish is SLOWER than the level loop (wave stage
            // 2.13 vs 1.96 ms): the lanes of a warp wait for each other in every one of the nested walks, so a warp still pays the
            // slowest ray of each phase.  What a small level costs is the latency of its longest ray (~80 us, grazing rays), either way.
            lp.pathTailMax = pathTailEnv >= 0 ? (uint32_t)pathTailEnv : 0u;
            BvhView v = view;
            unsigned long long* accData = acc->data;
            DeviceCounters* ctr = ctx->d_counters;
            void* args[] = {&v, &sp, &lp, &accData, &ctr};
            TPT_CUDA(cudaLaunchCooperativeKernel(kWaveLoop, dim3(ctx->wave_loop_grid[variant]), dim3(128), args, 0, st));
            mark();
            nLaunches += 1;
        }
        ctx->kernel_launches += nLaunches;
        TPT_CUDA(cudaGetLastError());
        if (timed) {
            TPT_CUDA(cudaEventSynchronize(ctx->stage_events[ev - 1]));
            for (size_t j = 1; j < ev; ++j) {
                float ms = 0.f;
                TPT_CUDA(cudaEventElapsedTime(&ms, ctx->stage_events[j - 1], ctx->stage_events[j]));
                const int stage = j == 1 ? 0 : 1;
                ctx->stage_ms[stage] += ms;
                ctx->stage_launches[stage] += 1;
                static const bool detail = getenv("TPT_STAGE_DETAIL") != nullptr;   // diagnostic: every launch's time on stderr
                if (detail) fprintf(stderr, "[tpt stage] batch@%u %s %.1f us\n", done, j == 1 ? "primary" : "wave", ms * 1e3f);
            }
        }
    }
    ctx->samples += (uint64_t)W * H * (uint64_t)cfg->spp;
    return TPT_OK;
}

int tpt_accum_add_impl(tpt_accum* dst, const tpt_accum* src) {
    const size_t n = (size_t)dst->width * dst->height * 3;
    k_accum_add<<<grid_for(n, 256, kNumSMs * 8), 256, 0, dst->ctx->stream>>>(dst->data, src->data, n);
    TPT_CUDA(cudaGetLastError());
    dst->ctx->kernel_launches += 1;
    return TPT_OK;
}

int tpt_accum_to_f32_impl(tpt_accum* acc, float* d_out) {
    const size_t n = (size_t)acc->width * acc->height * 3;
    k_accum_to_f32<<<grid_for(n, 256, kNumSMs * 8), 256, 0, acc->ctx->stream>>>(acc->data, d_out, n);
    TPT_CUDA(cudaGetLastError());
    acc->ctx->kernel_launches += 1;
    return TPT_OK;
}

int tpt_resolve_impl(tpt_accum* acc, int sppTotal, uint8_t* d_fb) {
    k_resolve<<<grid_for((size_t)acc->width * acc->height, 256, kNumSMs * 8), 256, 0, acc->ctx->stream>>>(
        acc->data, acc->width, acc->height, 1.0f / (float)sppTotal, reinterpret_cast<uchar4*>(d_fb));
    TPT_CUDA(cudaGetLastError());
    acc->ctx->kernel_launches += 1;
    return TPT_OK;
}

// ---- tpt_trace_rays_static: the batch trace entry for callers that must not allocate ----------------------------------
// No counterpart in the reference (include/tpt.h).  Same kernel as
// tpt_trace_rays (k_trace_rays_persistent), but every byte of device scratch it needs — the persistent kernel's work cursor
// and, for host buffers, the ray / hit staging — lives in ONE file-scope __device__ array of fixed size instead of
// cudaMalloc'd blocks.  Every GPU that loads the library holds the array (kStaticBytes, about 2.5 MiB) until exit; all calls
// on one GPU share it, so calls that overlap on the GPU race.  It keeps the last call's contents: the cursor is cleared by
// every call, the staging areas are overwritten before they are read.
namespace tpt {
constexpr size_t kStaticMaxRays = TPT_TRACE_STATIC_MAX_RAYS;
constexpr size_t kStaticRaysOff = 256;                                                   // [0, 256): the work cursor (8 bytes used)
constexpr size_t kStaticHitsOff = kStaticRaysOff + kStaticMaxRays * sizeof(tpt_ray);
constexpr size_t kStaticBytes = kStaticHitsOff + kStaticMaxRays * sizeof(tpt_hit);
static_assert(kStaticHitsOff % 16 == 0, "hits are written with 128-bit stores");
static __device__ __align__(256) unsigned char g_trace_static_scratch[kStaticBytes];
}   // namespace tpt

// In host code `g_trace_static_scratch` is only a handle for the symbol, not a device address: the address is looked up once per
// device, with that device current, and cached.
static int trace_static_scratch(int device, unsigned char** out) {
    constexpr int kMaxDevices = 64;
    static unsigned char* addr[kMaxDevices] = {};
    static std::mutex mu;
    TPT_REQUIRE(device >= 0 && device < kMaxDevices, "tpt_trace_rays_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (addr[device] == nullptr) {
        void* p = nullptr;
        TPT_CUDA(cudaGetSymbolAddress(&p, g_trace_static_scratch));   // the caller has made `device` current
        TPT_REQUIRE(p != nullptr, "tpt_trace_rays_static: cudaGetSymbolAddress returned NULL");
        addr[device] = static_cast<unsigned char*>(p);
    }
    *out = addr[device];
    return TPT_OK;
}

extern "C" int tpt_trace_rays_static(tpt_scene* s, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs) {
    TPT_REQUIRE(s && (n == 0 || (rays && hits)), "tpt_trace_rays_static: NULL argument");
    if (n > kStaticMaxRays) {
        set_error("tpt_trace_rays_static: " + std::to_string(n) + " rays exceed the static scratch (TPT_TRACE_STATIC_MAX_RAYS = " +
                  std::to_string(kStaticMaxRays) + "); split the batch or use tpt_trace_rays");
        return TPT_ERR_INVALID;
    }
    if (!s->built) {
        set_error("tpt_trace_rays_static: call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    if (n == 0) return TPT_OK;
    tpt_context* ctx = s->ctx;
    TPT_CUDA(cudaSetDevice(ctx->device));
    unsigned char* scratch = nullptr;
    TPT_CHECK(trace_static_scratch(ctx->device, &scratch));
    unsigned long long* cursor = reinterpret_cast<unsigned long long*>(scratch);
    const tpt_ray* dRays = rays;
    tpt_hit* dHits = hits;
    if (!device_ptrs) {
        dRays = reinterpret_cast<const tpt_ray*>(scratch + kStaticRaysOff);
        dHits = reinterpret_cast<tpt_hit*>(scratch + kStaticHitsOff);
        TPT_CUDA(cudaMemcpyAsync(scratch + kStaticRaysOff, rays, n * sizeof(tpt_ray), cudaMemcpyHostToDevice, ctx->stream));
    }
    TPT_CUDA(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long), ctx->stream));   // the array keeps the last call's cursor
    static unsigned grid[64] = {};   // an occupancy query, no allocation
    if (grid[ctx->device] == 0) grid[ctx->device] = kNumSMs * resident_blocks(k_trace_rays_persistent<false, false>, 128);
    k_trace_rays_persistent<false, false><<<grid[ctx->device], 128, 0, ctx->stream>>>(s->view(), dRays, n, dHits, any_hit, cursor, ctx->d_counters);
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += 1;
    if (!device_ptrs) {
        TPT_CUDA(cudaMemcpyAsync(hits, dHits, n * sizeof(tpt_hit), cudaMemcpyDeviceToHost, ctx->stream));
        TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return TPT_OK;
}
