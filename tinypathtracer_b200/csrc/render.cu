// render.cu — wavefront path tracer for sm_100a (generate / extend / shade / connect queues).
//
// Replaces the `trace` megakernel of the reference (src/path_tracer.cu:296-435), its ray generator
// sampleRays (:42-59), BSDF sampling getNewDirection (:187-225), delta-light NEE sampleDeltaLights
// (:265-286), env lookup sampleEnvLights (:288-294) and copyToFB (:451-471).
//
// Estimator.  The reference unwinds  radiance = (direct[d] + radiance) * atten[d] / p[d]  backwards
// over a per-thread stack.  With T_d = prod_{k<=d} atten_k / p_k this is, forwards,
//     L = sum_d T_d * direct_d + T_{D-1} * L_end,
// where atten/p == baseColor for all three lobes (diffuse: (|c|/pi) / (c/pi) with c > 0),
// direct_d = sum_lights visible * baseColor * Li  (+ emission of whatever the extra cosine "direct"
// ray hits, diffuse vertices only), and L_end is env(dir) on a miss, the grey emission when the
// vertex is emissive (which also discards that vertex's own direct term, :421-424), or 0 at the depth
// cut.  Every term is added to the pixel as soon as it is known, as an exact fixed-point integer, so
// the result does not depend on the order in which kernels, warps or GPUs finish.
#include "traverse.cuh"

namespace tpt {

// ---- path state: 3 x float4 per path ---------------------------------------------------------
//   s0 = (o.x, o.y, o.z, d.x)   s1 = (d.y, d.z, T.r, T.g)   s2 = (T.b, pixel, sample, -)
// secondary (shadow / direct) ray: same shape, T replaced by the contribution weight, s2.w = kind.
struct PathRec {
    float3 o, d, T;
    uint32_t pixel, sample;
};
__device__ __forceinline__ PathRec load_path(const float4* __restrict__ p, uint32_t i) {
    const float4 a = __ldcs(p + 3 * (size_t)i), b = __ldcs(p + 3 * (size_t)i + 1), c = __ldcs(p + 3 * (size_t)i + 2);
    PathRec r;
    r.o = f3(a.x, a.y, a.z);
    r.d = f3(a.w, b.x, b.y);
    r.T = f3(b.z, b.w, c.x);
    r.pixel = __float_as_uint(c.y);
    r.sample = __float_as_uint(c.z);
    return r;
}
__device__ __forceinline__ void store_path(float4* __restrict__ p, uint32_t i, float3 o, float3 d, float3 T, uint32_t pixel,
                                           uint32_t sample, uint32_t w) {
    __stcs(p + 3 * (size_t)i, make_float4(o.x, o.y, o.z, d.x));
    __stcs(p + 3 * (size_t)i + 1, make_float4(d.y, d.z, T.x, T.y));
    __stcs(p + 3 * (size_t)i + 2, make_float4(T.z, __uint_as_float(pixel), __uint_as_float(sample), __uint_as_float(w)));
}

// ---- accumulation: 3 x int64 fixed point per pixel ----------------------------------------------
__device__ __forceinline__ void accum_add(unsigned long long* __restrict__ acc, uint32_t pixel, float3 c, DeviceCounters* ctr) {
    if (!(isfinite(c.x) && isfinite(c.y) && isfinite(c.z))) {
        atomicAdd(&ctr->dropped_nonfinite, 1ull);
        return;
    }
    const float scale = 68719476736.0f;   // 2^36
    const float cap = 67108864.0f;        // 2^26 per contribution keeps int64 far from overflow
    unsigned long long* p = acc + 3 * (size_t)pixel;
    if (c.x != 0.f) atomicAdd(p + 0, (unsigned long long)__float2ll_rn(fminf(c.x, cap) * scale));
    if (c.y != 0.f) atomicAdd(p + 1, (unsigned long long)__float2ll_rn(fminf(c.y, cap) * scale));
    if (c.z != 0.f) atomicAdd(p + 2, (unsigned long long)__float2ll_rn(fminf(c.z, cap) * scale));
}

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
__device__ __forceinline__ void primary_ray(const CameraParams& cam, int px, int py, float ju, float jv, float3& o, float3& d) {
    // one rounding per operation, in the reference's order (bit-identical primary rays)
    float plx = xmul(xadd(ju, (float)px), cam.invW);
    float ply = xmul(xadd(jv, (float)py), cam.invH);
    plx = xmul(plx, cam.sensorW);
    ply = xmul(ply, cam.sensorH);
    const float3 dir = f3(xsub(plx, xmul(0.5f, cam.sensorW)), xsub(ply, xmul(0.5f, cam.sensorH)), -1.0f);
    d = xnormalize(xmat4_dir(cam.c2w, dir));
    o = xmat4_point(cam.c2w, f3(0.f, 0.f, 0.f));
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
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float* r = reinterpret_cast<const float*>(rays + i);
        const float2 a = __ldg(reinterpret_cast<const float2*>(r)), b = __ldg(reinterpret_cast<const float2*>(r) + 1),
                     c = __ldg(reinterpret_cast<const float2*>(r) + 2);
        const HitRec h = traverse<ANY, COUNT>(bvh, f3(a.x, a.y, b.x), f3(b.y, c.x, c.y), &ctr->node_visits, &ctr->tri_tests);
        reinterpret_cast<float4*>(hits)[i] = make_float4(__int_as_float(h.prim), h.t, h.u, h.v);
    }
}

// ---- generate ------------------------------------------------------------------------------------
// Batch = all pixels x `k` consecutive samples.  Slot i -> (tile-ordered pixel, local sample).
__global__ void __launch_bounds__(256) k_generate(CameraParams cam, int W, int H, uint32_t nPadded, uint32_t kSamples, uint32_t sampleBase,
                                                  uint32_t sampleStride, uint64_t seedKey, float4* __restrict__ paths,
                                                  uint32_t* __restrict__ counts) {
    const uint32_t total = nPadded * kSamples;
    for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += gridDim.x * blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        bool valid = i < total;
        int px = 0, py = 0;
        uint32_t sIdx = 0;
        if (valid) {
            const uint32_t sLocal = i / nPadded, p = i - sLocal * nPadded;
            valid = tile_decode(p, W, H, px, py);
            sIdx = sampleBase + sLocal * sampleStride;
        }
        const uint32_t slot = warp_append(valid, counts + 0);
        if (valid) {
            const uint32_t pixel = (uint32_t)(px + py * W);
            float ju, jv;
            rng_pair(rng_path_key(seedKey, pixel, sIdx), 0u, ju, jv);
            float3 o, d;
            primary_ray(cam, px, py, ju, jv, o, d);
            store_path(paths, slot, o, d, f3(1.f, 1.f, 1.f), pixel, sIdx, 0u);
        }
    }
}

// ---- extend: closest hit for every live path --------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(128) k_extend(BvhView bvh, const float4* __restrict__ paths, const uint32_t* __restrict__ countIn,
                                                uint32_t* __restrict__ countNext, uint32_t* __restrict__ countSec, float4* __restrict__ hits,
                                                DeviceCounters* ctr, int depth) {
    const uint32_t n = *countIn;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // The shade kernel of this bounce appends to these; nothing reads them before it runs.
        *countNext = 0u;
        *countSec = 0u;
        atomicAdd(depth == 0 ? &ctr->rays_primary : &ctr->rays_extension, (unsigned long long)n);
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = __ldg(paths + 3 * (size_t)i), b = __ldg(paths + 3 * (size_t)i + 1);
        const HitRec h = traverse<false, COUNT>(bvh, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), &ctr->node_visits, &ctr->tri_tests);
        __stcs(hits + i, make_float4(h.t, h.u, h.v, __int_as_float(h.leaf)));
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
// clamp addressing, normalised coordinates (src/texture.cu:160-168).
__device__ __forceinline__ float3 env_lookup(const uchar4* __restrict__ env, int ew, int eh, float3 d) {
    if (env == nullptr) return f3(0.f, 0.f, 0.f);
    float u = atan2f(d.z, d.x) / (2.0f * kPi);
    if (u < 0.0f) u += 1.0f;
    const float v = 1.0f - acosf(fmaxf(fminf(d.y, 1.0f), -1.0f)) / kPi;
    int x = (int)floorf(u * (float)ew), y = (int)floorf(v * (float)eh);
    x = min(max(x, 0), ew - 1);
    y = min(max(y, 0), eh - 1);
    const uchar4 t = __ldg(env + (size_t)y * ew + x);
    return f3((float)t.x, (float)t.y, (float)t.z) * (1.0f / 255.0f);
}

struct ShadeParams {
    const float4* mtl;       // 2 float4 per material
    const tpt_delta_light* lights;
    int nLights;
    const uchar4* env;
    int envW, envH;
    int depth, maxDepth;
    uint64_t seedKey;
};

// ---- shade: material evaluation, NEE ray emission, path continuation ------------------------------
__global__ void __launch_bounds__(256) k_shade(BvhView bvh, ShadeParams sp, const float4* __restrict__ pathsIn, const float4* __restrict__ hits,
                                               const uint32_t* __restrict__ countIn, float4* __restrict__ pathsOut, uint32_t* __restrict__ countNext,
                                               float4* __restrict__ sec, uint32_t* __restrict__ countSec, unsigned long long* __restrict__ acc,
                                               DeviceCounters* ctr) {
    const uint32_t n = *countIn;
    // every lane of a warp must reach the warp_append calls: iterate whole warps
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const bool live = i < n;
        bool wantNext = false, wantDirect = false;
        int nShadow = 0;
        PathRec p{};
        float3 P = f3(0, 0, 0), nrm = f3(0, 0, 0), base3 = f3(0, 0, 0), Tn = f3(0, 0, 0);
        float eta = 0.f, metallic = 0.f;
        uint64_t key = 0;
        if (live) {
            p = load_path(pathsIn, i);
            const float4 h = __ldcs(hits + i);
            const int leaf = __float_as_int(h.w);
            if (leaf < 0) {
                accum_add(acc, p.pixel, p.T * env_lookup(sp.env, sp.envW, sp.envH, p.d), ctr);
            } else {
                const float4 g2 = __ldg(bvh.tri_geom + 3 * (size_t)leaf + 2);
                const int mtlIdx = __float_as_int(g2.z);
                const float4 m0 = __ldg(sp.mtl + 2 * mtlIdx), m1 = __ldg(sp.mtl + 2 * mtlIdx + 1);
                if (m0.w > 0.0f) {
                    accum_add(acc, p.pixel, p.T * m0.w, ctr);   // emissive vertex: terminal, grey emission
                } else {
                    const float4 na = __ldg(bvh.tri_norm + 3 * (size_t)leaf), nb = __ldg(bvh.tri_norm + 3 * (size_t)leaf + 1),
                                 nc = __ldg(bvh.tri_norm + 3 * (size_t)leaf + 2);
                    const float u = h.y, v = h.z, w = xsub(xsub(1.0f, u), v);
                    const float3 n0 = f3(na.x, na.y, na.z), n1 = f3(na.w, nb.x, nb.y), n2 = f3(nb.z, nb.w, nc.x);
                    nrm = xnormalize(f3(xadd(xadd(xmul(w, n0.x), xmul(u, n1.x)), xmul(v, n2.x)),
                                        xadd(xadd(xmul(w, n0.y), xmul(u, n1.y)), xmul(v, n2.y)),
                                        xadd(xadd(xmul(w, n0.z), xmul(u, n1.z)), xmul(v, n2.z))));
                    P = f3(xadd(p.o.x, xmul(h.x, p.d.x)), xadd(p.o.y, xmul(h.x, p.d.y)), xadd(p.o.z, xmul(h.x, p.d.z)));
                    base3 = f3(m0.x, m0.y, m0.z);
                    eta = m1.x;
                    metallic = m1.y;
                    Tn = p.T * base3;
                    key = rng_path_key(sp.seedKey, p.pixel, p.sample);
                    nShadow = sp.nLights;
                    wantDirect = !(eta >= 1.0f || metallic > 0.0f);
                    wantNext = sp.depth + 1 < sp.maxDepth;
                }
            }
        }
        // delta lights: one infinite shadow ray per light, contribution T_d * baseColor * Li
        for (int l = 0; l < sp.nLights; ++l) {
            const bool emit = l < nShadow;
            float3 ldir = f3(0, 0, 1), lrad = f3(0, 0, 0);
            if (emit) light_incoming(sp.lights[l], P, ldir, lrad);
            const uint32_t slot = warp_append(emit, countSec);
            if (emit) store_path(sec, slot, P, ldir, Tn * (base3 * lrad), p.pixel, p.sample, 0u);
        }
        // the reference's extra cosine "direct" ray (diffuse only): adds emission of what it hits
        {
            float3 ddir = f3(0, 0, 1);
            if (wantDirect) {
                float a, b;
                rng_pair(key, 2u + 2u * (uint32_t)sp.depth, a, b);
                wantDirect = new_direction(p.d, nrm, eta, metallic, a, b, ddir);
                if (!wantDirect) atomicAdd(&ctr->dropped_nonfinite, 1ull);
            }
            const uint32_t slot = warp_append(wantDirect, countSec);
            if (wantDirect) store_path(sec, slot, P, ddir, Tn, p.pixel, p.sample, 1u);
        }
        {
            float3 ndir = f3(0, 0, 1);
            if (wantNext) {
                float a, b;
                rng_pair(key, 1u + 2u * (uint32_t)sp.depth, a, b);
                wantNext = new_direction(p.d, nrm, eta, metallic, a, b, ndir);
                if (!wantNext) atomicAdd(&ctr->dropped_nonfinite, 1ull);
                wantNext = wantNext && (Tn.x > 0.f || Tn.y > 0.f || Tn.z > 0.f);   // a black path adds nothing further
            }
            const uint32_t slot = warp_append(wantNext, countNext);
            if (wantNext) store_path(pathsOut, slot, P, ndir, Tn, p.pixel, p.sample, 0u);
        }
    }
}

// ---- connect: shadow rays (any hit, infinite) and direct rays (closest hit -> emission) --------------
template <bool COUNT>
__global__ void __launch_bounds__(128) k_connect(BvhView bvh, const float4* __restrict__ mtl, const float4* __restrict__ sec,
                                                 const uint32_t* __restrict__ countSec, unsigned long long* __restrict__ acc, DeviceCounters* ctr) {
    const uint32_t n = *countSec;
    unsigned nDirect = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = __ldcs(sec + 3 * (size_t)i), b = __ldcs(sec + 3 * (size_t)i + 1), c = __ldcs(sec + 3 * (size_t)i + 2);
        const float3 o = f3(a.x, a.y, a.z), d = f3(a.w, b.x, b.y), wgt = f3(b.z, b.w, c.x);
        const uint32_t pixel = __float_as_uint(c.y), kind = __float_as_uint(c.w);
        if (kind == 0u) {
            const HitRec h = traverse<true, COUNT>(bvh, o, d, &ctr->node_visits, &ctr->tri_tests);
            if (h.leaf < 0) accum_add(acc, pixel, wgt, ctr);
        } else {
            ++nDirect;
            const HitRec h = traverse<false, COUNT>(bvh, o, d, &ctr->node_visits, &ctr->tri_tests);
            if (h.leaf >= 0) {
                const int mtlIdx = __float_as_int(__ldg(bvh.tri_geom + 3 * (size_t)h.leaf + 2).z);
                const float e = __ldg(mtl + 2 * mtlIdx).w;
                if (e != 0.0f) accum_add(acc, pixel, wgt * e, ctr);
            }
        }
    }
    // ray accounting: one atomic per block
    __shared__ unsigned sDirect;
    if (threadIdx.x == 0) sDirect = 0;
    __syncthreads();
    if (nDirect) atomicAdd(&sDirect, nDirect);
    __syncthreads();
    if (threadIdx.x == 0) {
        if (sDirect) atomicAdd(&ctr->rays_direct, (unsigned long long)sDirect);
        if (blockIdx.x == 0) atomicAdd(&ctr->rays_shadow, (unsigned long long)n);   // = shadow + direct; host subtracts
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

// copyToFB, src/path_tracer.cu:451-471 + Spectrum::toUChar, include/material.h:74-81:
// (unsigned char) clamp(255 * L / spp, 0, 255) by truncation, B,G,R order, alpha untouched, row H-1-y.
__device__ __forceinline__ unsigned char to_uchar(float v) {
    const float x = v * 255.0f;
    const float c = fmaxf((x < 255.0f) ? x : 255.0f, 0.0f);   // min(x,255) with the reference's NaN -> 255
    return (unsigned char)c;
}
__global__ void __launch_bounds__(256) k_resolve(const unsigned long long* __restrict__ acc, int W, int H, float invSpp, uchar4* __restrict__ fb) {
    const uint32_t n = (uint32_t)W * (uint32_t)H;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int x = (int)(i % (uint32_t)W), y = (int)(i / (uint32_t)W);
        const unsigned long long* p = acc + 3 * (size_t)i;
        const float r = (float)((double)(long long)p[0] * (1.0 / 68719476736.0)) * invSpp;
        const float g = (float)((double)(long long)p[1] * (1.0 / 68719476736.0)) * invSpp;
        const float b = (float)((double)(long long)p[2] * (1.0 / 68719476736.0)) * invSpp;
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

int tpt_primary_rays_impl(tpt_context* ctx, const tpt_camera* cam, int W, int H, int jitter, uint64_t seed, tpt_ray* d_rays) {
    k_primary_rays<<<grid_for((size_t)W * H, 256, kNumSMs * 8), 256, 0, ctx->stream>>>(make_camera(cam, W, H), W, H, jitter, rng_seed_key(seed), d_rays);
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += 1;
    return TPT_OK;
}

int tpt_trace_rays_impl(tpt_scene* s, const tpt_ray* d_rays, size_t n, tpt_hit* d_hits, int anyHit, int counted) {
    tpt_context* ctx = s->ctx;
    const unsigned grid = grid_for(n, 128, kNumSMs * 16);
    const BvhView v = s->view();
    if (anyHit) {
        if (counted) k_trace_rays<true, true><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
        else k_trace_rays<true, false><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
    } else {
        if (counted) k_trace_rays<false, true><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
        else k_trace_rays<false, false><<<grid, 128, 0, ctx->stream>>>(v, d_rays, n, d_hits, ctx->d_counters);
    }
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += 1;
    return TPT_OK;
}

static int ensure_pools(tpt_context* ctx, size_t paths, size_t secPerPath) {
    const size_t secNeed = paths * secPerPath;
    if (paths > ctx->pool_paths) {
        TPT_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 2; ++k) {
            if (ctx->path_state[k]) cudaFree(ctx->path_state[k]);
            TPT_CUDA(cudaMalloc(&ctx->path_state[k], paths * 3 * sizeof(float4)));
        }
        if (ctx->hits) cudaFree(ctx->hits);
        TPT_CUDA(cudaMalloc(&ctx->hits, paths * sizeof(float4)));
        ctx->pool_paths = paths;
    }
    if (secNeed > ctx->sec_capacity) {
        TPT_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->sec_rays) cudaFree(ctx->sec_rays);
        TPT_CUDA(cudaMalloc(&ctx->sec_rays, secNeed * 3 * sizeof(float4)));
        ctx->sec_capacity = secNeed;
    }
    return TPT_OK;
}

int tpt_render_impl(tpt_scene* s, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc) {
    tpt_context* ctx = s->ctx;
    cudaStream_t st = ctx->stream;
    const int W = cfg->width, H = cfg->height;
    const uint32_t nPadded = tiles_x(W) * tiles_y(H) * 32u;
    size_t target = cfg->paths_in_flight > 0 ? (size_t)cfg->paths_in_flight : (size_t)1 << 21;
    uint32_t k = (uint32_t)(target / nPadded);
    if (k < 1) k = 1;
    if (k > (uint32_t)cfg->spp) k = (uint32_t)cfg->spp;
    const size_t pool = (size_t)nPadded * k;
    TPT_CHECK(ensure_pools(ctx, pool, s->n_lights + 1));

    const CameraParams cp = make_camera(cam, W, H);
    const uint64_t seedKey = rng_seed_key(cfg->seed);
    const bool counted = (cfg->flags & TPT_RENDER_COUNTERS) != 0;
    const bool timed = (cfg->flags & TPT_RENDER_STAGE_TIMES) != 0;
    const size_t nEvents = 2 + 3 * (size_t)cfg->max_depth;
    if (timed) {
        while (ctx->stage_events.size() < nEvents) {
            cudaEvent_t e;
            TPT_CUDA(cudaEventCreate(&e));
            ctx->stage_events.push_back(e);
        }
    }
    size_t ev = 0;
    auto mark = [&]() { if (timed) cudaEventRecord(ctx->stage_events[ev++], st); };
    const BvhView view = s->view();
    uint32_t* counts = ctx->queue_counts;
    const unsigned gridGen = kNumSMs * 8, gridTrace = kNumSMs * 16, gridShade = kNumSMs * 8;

    ShadeParams sp;
    sp.mtl = s->mtl_table;
    sp.lights = s->lights;
    sp.nLights = (int)s->n_lights;
    sp.env = ctx->env;
    sp.envW = ctx->env_w;
    sp.envH = ctx->env_h;
    sp.maxDepth = cfg->max_depth;
    sp.seedKey = seedKey;

    for (uint32_t done = 0; done < (uint32_t)cfg->spp; done += k) {
        const uint32_t kk = ((uint32_t)cfg->spp - done) < k ? ((uint32_t)cfg->spp - done) : k;
        TPT_CUDA(cudaMemsetAsync(counts, 0, 4 * sizeof(uint32_t), st));
        ev = 0;
        mark();
        k_generate<<<gridGen, 256, 0, st>>>(cp, W, H, nPadded, kk, (uint32_t)cfg->sample_offset + done * (uint32_t)cfg->sample_stride,
                                            (uint32_t)cfg->sample_stride, seedKey, ctx->path_state[0], counts);
        ctx->kernel_launches += 1;
        mark();
        for (int depth = 0; depth < cfg->max_depth; ++depth) {
            const int cur = depth & 1, nxt = cur ^ 1;
            sp.depth = depth;
            if (counted)
                k_extend<true><<<gridTrace, 128, 0, st>>>(view, ctx->path_state[cur], counts + cur, counts + nxt, counts + 2, ctx->hits,
                                                          ctx->d_counters, depth);
            else
                k_extend<false><<<gridTrace, 128, 0, st>>>(view, ctx->path_state[cur], counts + cur, counts + nxt, counts + 2, ctx->hits,
                                                           ctx->d_counters, depth);
            mark();
            k_shade<<<gridShade, 256, 0, st>>>(view, sp, ctx->path_state[cur], ctx->hits, counts + cur, ctx->path_state[nxt], counts + nxt,
                                               ctx->sec_rays, counts + 2, acc->data, ctx->d_counters);
            mark();
            if (counted) k_connect<true><<<gridTrace, 128, 0, st>>>(view, s->mtl_table, ctx->sec_rays, counts + 2, acc->data, ctx->d_counters);
            else k_connect<false><<<gridTrace, 128, 0, st>>>(view, s->mtl_table, ctx->sec_rays, counts + 2, acc->data, ctx->d_counters);
            mark();
            ctx->kernel_launches += 3;
        }
        TPT_CUDA(cudaGetLastError());
        if (timed) {
            TPT_CUDA(cudaEventSynchronize(ctx->stage_events[ev - 1]));
            for (size_t k = 1; k < ev; ++k) {
                float ms = 0.f;
                TPT_CUDA(cudaEventElapsedTime(&ms, ctx->stage_events[k - 1], ctx->stage_events[k]));
                const int stage = k == 1 ? 0 : 1 + (int)((k - 2) % 3);
                ctx->stage_ms[stage] += ms;
                ctx->stage_launches[stage] += 1;
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
