// oracle.cpp — CPU restatement of the TinyPathTracer hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under tinypathtracer_b200/ may include, link or call this
// file; it is used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs as the checker and the CPU baseline.
//
// Every function cites the reference file:line it restates (paths relative to the reference
// repository root).  All arithmetic is fp32, written in the reference's operand order, and this
// file is compiled with -ffp-contract=off so no multiply-add is fused: that is what the reference's
// CUDA_CALLABLE code evaluates to when compiled for the host (x86-64 without FMA contraction).
//
// Parity status: the reference repository has no tests or golden vectors ("parity unpinned" by the
// reference itself).  This restatement is pinned instead against the reference's own host-callable
// functions compiled from /root/reference (oracle/ref_host.cu -> oracle/_ref/libref_host.so) and the
// golden vectors that harness produced (tests/golden/ref_host_vectors.npz).  The RNG is NOT the
// reference's cuRAND XORWOW (seeded by time(), irreproducible): both this oracle and the CUDA build
// use the counter-based generator specified in DESIGN.md, so RNG-dependent quantities agree with the
// reference only in distribution.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#if defined(_OPENMP)
#include <omp.h>
#endif

namespace {

// ----------------------------------------------------------------------------------------------
// include/math/vec.h
// ----------------------------------------------------------------------------------------------
constexpr float kPi = 3.141592653589793f;   // vec.h:66
constexpr float kDelta = 2e-4f;             // vec.h:70

struct V3 {
    float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }   // vec.h:137
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }   // vec.h:138
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }                        // vec.h:139
inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }      // vec.h:140
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }      // vec.h:141
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }   // vec.h:142
inline V3 operator/(V3 a, float s) { return (1.0f / s) * a; }                     // vec.h:143
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }        // vec.h:151
inline V3 cross(V3 a, V3 b) {                                                    // vec.h:152
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float norm2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }            // vec.h:130
inline float norm(V3 a) { return sqrtf(norm2(a)); }                               // vec.h:131
inline float tmin(float x, float y) { return x < y ? x : y; }                     // vec.h:76
inline float tmax(float x, float y) { return x > y ? x : y; }                     // vec.h:74
inline float clampf(float x, float ub, float lb) { return tmax(tmin(x, ub), lb); }  // vec.h:78
inline float saturate(float x) { return clampf(x, 1.0f, 0.0f); }                  // vec.h:82
inline float square(float x) { return x * x; }                                    // vec.h:80
inline V3 hmin(V3 a, V3 b) { return v3(tmin(a.x, b.x), tmin(a.y, b.y), tmin(a.z, b.z)); }  // vec.h:154
inline V3 hmax(V3 a, V3 b) { return v3(tmax(a.x, b.x), tmax(a.y, b.y), tmax(a.z, b.z)); }  // vec.h:153

// vec.h:25-38 — Quake reciprocal square root, one Newton step.
inline float frsqrt(float num) {
    int32_t i;
    float x2 = num * 0.5f, y = num;
    std::memcpy(&i, &y, 4);
    i = 0x5f3759df - (i >> 1);
    std::memcpy(&y, &i, 4);
    y = y * (1.5f - (x2 * y * y));
    return y;
}
inline V3 normalize(V3 v) { return v * frsqrt(norm2(v)); }                        // vec.h:155

// include/math/mat.h:37-51 — column-major Mat4 times Vec4, accumulation order res[i] += m[j][i]*v[j].
inline void mat4_mul_vec4(const float* m, const float v[4], float out[4]) {
    for (int i = 0; i < 4; ++i) {
        float r = 0.0f;
        for (int j = 0; j < 4; ++j) r += m[j * 4 + i] * v[j];
        out[i] = r;
    }
}
inline V3 xform_point(const float* m, V3 p) {
    float v[4] = {p.x, p.y, p.z, 1.0f}, o[4];
    mat4_mul_vec4(m, v, o);
    return v3(o[0], o[1], o[2]);
}
inline V3 xform_dir(const float* m, V3 d) {
    float v[4] = {d.x, d.y, d.z, 0.0f}, o[4];
    mat4_mul_vec4(m, v, o);
    return v3(o[0], o[1], o[2]);
}
// mat.h:18-35 — MatrixMultiply, res[j][i] += lhs[k][i] * rhs[j][k].
inline void mat4_mul(const float* a, const float* b, float* out) {
    float r[16];
    for (int i = 0; i < 16; ++i) r[i] = 0.0f;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            for (int k = 0; k < 4; ++k) r[j * 4 + i] += a[k * 4 + i] * b[j * 4 + k];
    std::memcpy(out, r, sizeof r);
}

// ----------------------------------------------------------------------------------------------
// Scene views (the DeviceScene contract, include/mesh.cuh:80-96)
// ----------------------------------------------------------------------------------------------
struct Material {   // include/material.h:86-120, 60 bytes
    float base[3];
    float emission, eta, metallic;
    float unused[9];
};
static_assert(sizeof(Material) == 60, "Material layout");
struct MtlInterval {   // include/mesh.cuh:74-78
    int32_t begin, mtl;
};
struct DeltaLight {   // include/delta_light.h:96-130, 52 bytes: type + union{point,directional,spot}
    int32_t type;
    float color[3];
    float intensity;
    float a[3];   // point/spot: pos ; directional: direction
    float b[3];   // spot: direction
    float cosOuter, invConeDiff;
};
static_assert(sizeof(DeltaLight) == 52, "DeltaLight layout");

struct RefNode {   // include/bvh.cuh:34-58, 36 bytes
    uint32_t parent;
    int32_t a, b;   // intern{left,right} | leaf{fid,pad}
    float mn[3], mx[3];
};
static_assert(sizeof(RefNode) == 36, "BVHNode layout");

struct Hit {   // src/path_tracer.cu:19-26
    int32_t idx;
    float dist, u, v;
};
inline Hit no_hit() { return Hit{-1, FLT_MAX, 0.0f, 0.0f}; }

struct Ray {   // include/ray.h:8-16
    V3 o, d;
};

inline V3 ld3(const float* p, size_t i) { return v3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }

// ----------------------------------------------------------------------------------------------
// include/geometry_queries.h
// ----------------------------------------------------------------------------------------------
// geometry_queries.h:18-46.  `max`/`min` on floats in a .cu translation unit resolve to CUDA's
// non-template overloads (fmaxf/fminf: a NaN operand yields the other operand); comparisons are
// plain IEEE and false on NaN.
inline bool ray_hit_bbox(const Ray& r, const float mn[3], const float mx[3]) {
    float inv[3] = {1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z};
    float org[3] = {r.o.x, r.o.y, r.o.z};
    float t0 = -FLT_MAX, t1 = FLT_MAX;
    for (int a = 0; a < 3; ++a) {
        float lo = (mn[a] - org[a]) * inv[a];
        float hi = (mx[a] - org[a]) * inv[a];
        if (lo > hi) std::swap(lo, hi);
        if (t0 > hi || lo > t1) return false;
        t0 = fmaxf(t0, lo);
        t1 = fminf(t1, hi);
    }
    return true;
}

// geometry_queries.h:66-86 — Möller–Trumbore, two-sided, inclusive edges.
inline bool ray_hit_triangle(const Ray& r, V3 v0, V3 v1, V3 v2, float& dist, float& u_out, float& v_out) {
    V3 e1 = v1 - v0;
    V3 e2 = v2 - v0;
    V3 t = r.o - v0;
    V3 p = cross(r.d, e2);
    V3 q = cross(t, e1);
    float denom = dot(p, e1);
    if (denom == 0.0f) return false;
    float invDenom = 1.0f / denom;
    float u = dot(p, t) * invDenom;
    float v = dot(q, r.d) * invDenom;
    if (u < 0.0f || v < 0.0f || u + v > 1.0f) return false;
    u_out = u;
    v_out = v;
    dist = dot(q, e2) * invDenom;
    return true;
}

// ----------------------------------------------------------------------------------------------
// src/bvh.cu — Morton keys, hierarchy, refit
// ----------------------------------------------------------------------------------------------
inline void expand_bits(int64_t& bits) {   // bvh.cu:14-21
    bits = (bits | bits << 32) & 0x1f00000000ffff;
    bits = (bits | bits << 16) & 0x1f0000ff0000ff;
    bits = (bits | bits << 8) & 0x100f00f00f00f00f;
    bits = (bits | bits << 4) & 0x10c30c30c30c30c3;
    bits = (bits | bits << 2) & 0x1249249249249249;
}
inline int64_t float_to_21(float x) {   // bvh.cu:23-46
    int32_t ix;
    std::memcpy(&ix, &x, 4);
    int32_t exponent = (ix >> 23) & 0xff;
    int32_t mantissa = (ix & 0x00ffffff) | 0x00800000;
    exponent -= 127;
    uint32_t signBit = (uint32_t)ix & 0x80000000u;
    signBit >>= 30;
    int32_t sign = -1 * ((int32_t)signBit - 1);
    int32_t value;
    if (exponent >= 8) {
        value = INT32_MAX;
    } else if (exponent >= 0) {
        value = (int32_t)((uint32_t)mantissa << exponent);
    } else {
        // The reference shifts by -exponent unguarded; a shift count >= 32 is what the GPU clamps
        // to 0.  Restated with the GPU's behaviour.
        int sh = -exponent;
        value = sh >= 32 ? 0 : (mantissa >> sh);
    }
    value = (int32_t)((int64_t)value * sign);
    value = (int32_t)((uint32_t)value + (uint32_t)INT32_MAX);
    uint32_t bits = (uint32_t)value;
    return (int64_t)((bits & 0xfffff800u) >> 11);
}
inline int64_t morton_code(V3 p) {   // bvh.cu:50-62
    int64_t ix = float_to_21(p.x), iy = float_to_21(p.y), iz = float_to_21(p.z);
    expand_bits(ix);
    expand_bits(iy);
    expand_bits(iz);
    return ix | (iy << 1) | (iz << 2);
}
inline int clzll(int64_t x) { return x == 0 ? 64 : __builtin_clzll((unsigned long long)x); }   // bvh.cu:9-12

// bvh.cu:64-99
int get_the_other_end(const int64_t* keys, int idx, int size, int& dir, int& lmax) {
    int64_t left = idx == 0 ? -1 : keys[idx - 1];
    int64_t right = keys[idx + 1];
    int64_t self = keys[idx];
    int leftCommon = clzll(left ^ self);
    int rightCommon = clzll(right ^ self);
    dir = leftCommon > rightCommon ? -1 : 1;
    int minRange = std::min(leftCommon, rightCommon);
    lmax = 2;
    int other = idx + dir * lmax;
    while (other >= 0 && other < size && clzll(self ^ keys[other]) > minRange) {
        lmax <<= 1;
        other = idx + dir * lmax;
    }
    int range = 0;
    for (int step = lmax >> 1; step > 0; step >>= 1) {
        other = idx + (range + step) * dir;
        if (other < 0 || other >= size) continue;
        if (clzll(self ^ keys[other]) > minRange) range += step;
    }
    return idx + range * dir;
}
// bvh.cu:101-120
int find_split_position(const int64_t* keys, int idx, int otherEnd, int dir, int lmax) {
    int left, right;
    if (dir == -1) {
        left = otherEnd;
        right = idx;
    } else {
        left = idx;
        right = otherEnd;
    }
    int delta = clzll(keys[left] ^ keys[right]);
    int split = 0;
    for (int t = lmax >> 1; t > 0; t >>= 1) {
        int pos = idx + dir * (split + t);
        if (pos < left || pos > right) continue;
        if (clzll(keys[idx] ^ keys[pos]) > delta) split += t;
    }
    return idx + split * dir;
}

// src/bvh.cu:304-331 BVH::construct = initNodes (:128-148) + stable sort by key (:326, thrust radix
// sort) + computeNodeRange (:150-217) + computeBBox (:220-302).  Returns 0, or 1 if the level
// iteration of computeBBox would not terminate (malformed hierarchy).
int build_reference_bvh(const float* wverts, const uint32_t* indices, int F, RefNode* nodes, int64_t* keys) {
    RefNode* iN = nodes;
    RefNode* lN = nodes + (F - 1);
    for (int i = 0; i < 2 * F - 1; ++i) {
        nodes[i].parent = 0;
        nodes[i].a = nodes[i].b = 0;
        for (int k = 0; k < 3; ++k) {
            nodes[i].mn[k] = FLT_MAX;
            nodes[i].mx[k] = -FLT_MAX;
        }
    }
    std::vector<int64_t> k0(F);
    std::vector<RefNode> l0(F);
    for (int f = 0; f < F; ++f) {   // initNodes
        V3 a = ld3(wverts, indices[3 * f]), b = ld3(wverts, indices[3 * f + 1]), c = ld3(wverts, indices[3 * f + 2]);
        V3 mn = hmin(hmin(a, b), c), mx = hmax(hmax(a, b), c);
        V3 ctr = (mn + mx) * 0.5f;
        k0[f] = morton_code(ctr);
        RefNode n{};
        n.parent = 0;
        n.a = f;
        n.b = 0;
        n.mn[0] = mn.x; n.mn[1] = mn.y; n.mn[2] = mn.z;
        n.mx[0] = mx.x; n.mx[1] = mx.y; n.mx[2] = mx.z;
        l0[f] = n;
    }
    std::vector<int> order(F);
    for (int i = 0; i < F; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return k0[x] < k0[y]; });
    for (int i = 0; i < F; ++i) {
        keys[i] = k0[order[i]];
        lN[i] = l0[order[i]];
    }
    const int nFaces = F;
    for (int t = 0; t < nFaces - 1; ++t) {   // computeNodeRange
        int dir = 0, lmax = 0;
        int otherEnd = get_the_other_end(keys, t, nFaces, dir, lmax);
        int splitPos = find_split_position(keys, t, otherEnd, dir, lmax);
        if (dir == 1) {
            if (std::max(t, otherEnd) == splitPos + 1) {
                lN[splitPos + 1].parent = t;
                iN[t].b = splitPos + nFaces;
            } else {
                iN[splitPos + 1].parent = t;
                iN[t].b = splitPos + 1;
            }
            if (std::min(t, otherEnd) == splitPos) {
                lN[splitPos].parent = t;
                iN[t].a = splitPos + nFaces - 1;
            } else {
                iN[splitPos].parent = t;
                iN[t].a = splitPos;
            }
        } else {
            if (std::min(t, otherEnd) == splitPos - 1) {
                lN[splitPos - 1].parent = t;
                iN[t].a = splitPos + nFaces - 2;
            } else {
                iN[splitPos - 1].parent = t;
                iN[t].a = splitPos - 1;
            }
            if (std::max(t, otherEnd) == splitPos) {
                lN[splitPos].parent = t;
                iN[t].b = splitPos + nFaces - 1;
            } else {
                iN[splitPos].parent = t;
                iN[t].b = splitPos;
            }
        }
    }
    // computeBBox: levels by fixed-point iteration, then level-by-level union.
    const int interSize = F - 1;
    if (interSize <= 0) return 0;
    std::vector<int> level(interSize, INT32_MAX);
    for (int i = 0; i < interSize; ++i)
        if (iN[i].a >= interSize && iN[i].b >= interSize) level[i] = 0;
    int currentLevel = 0;
    bool flag = true;
    // The single-leaf-pair root case: reference spins while(flag) until node 0 gets a level through
    // `cond`; with interSize == 1 that never happens.  Bounded here.
    int guard = 0;
    while (flag) {
        for (int i = 0; i < interSize; ++i) {
            int l = iN[i].a, r = iN[i].b;
            bool cond = (l < interSize && level[l] == currentLevel && ((r < interSize && level[r] <= currentLevel) || r >= interSize)) ||
                        (r < interSize && level[r] == currentLevel && ((l < interSize && level[l] <= currentLevel) || l >= interSize));
            if (cond) {
                level[i] = currentLevel + 1;
                if (i == 0) flag = false;
            }
        }
        currentLevel += 1;
        if (++guard > 2 * F + 64) {
            if (interSize == 1) break;   // F == 2: root is level 0 already
            return 1;
        }
    }
    for (int lv = 0; lv <= currentLevel; ++lv) {
        for (int i = 0; i < interSize; ++i) {
            if (level[i] != lv) continue;
            const RefNode& L = nodes[iN[i].a];
            const RefNode& R = nodes[iN[i].b];
            for (int k = 0; k < 3; ++k) {
                iN[i].mn[k] = tmin(L.mn[k], R.mn[k]);
                iN[i].mx[k] = tmax(L.mx[k], R.mx[k]);
            }
        }
    }
    return 0;
}

// src/path_tracer.cu:61-107 — unordered stack traversal, child boxes tested, first-met wins ties.
// Returns false if the 64-entry stack of the reference would overflow (UB there).
bool traverse_reference(const Ray& ray, const RefNode* nodes, const float* wverts, const uint32_t* indices, int nFaces,
                        Hit& status, uint64_t* nInt = nullptr, uint64_t* nLeaf = nullptr) {
    if (nFaces == 1) {
        // Degenerate: node 0 is the only leaf (0 >= nFaces-1).
    }
    int stack[64];
    int sp = 1;
    stack[0] = 0;
    bool ok = true;
    while (sp > 0) {
        int cur = stack[--sp];
        if (cur >= nFaces - 1) {
            int fid = nodes[cur].a;
            float dist, u, v;
            if (nLeaf) ++*nLeaf;
            if (ray_hit_triangle(ray, ld3(wverts, indices[3 * fid]), ld3(wverts, indices[3 * fid + 1]),
                                 ld3(wverts, indices[3 * fid + 2]), dist, u, v)) {
                if (dist < status.dist && dist > kDelta) {
                    status.dist = dist;
                    status.idx = fid;
                    status.u = u;
                    status.v = v;
                }
            }
        } else {
            int l = nodes[cur].a, r = nodes[cur].b;
            if (nInt) ++*nInt;
            if (ray_hit_bbox(ray, nodes[l].mn, nodes[l].mx)) {
                if (sp >= 64) { ok = false; break; }
                stack[sp++] = l;
            }
            if (ray_hit_bbox(ray, nodes[r].mn, nodes[r].mx)) {
                if (sp >= 64) { ok = false; break; }
                stack[sp++] = r;
            }
        }
    }
    return ok;
}

// Authority for closest hit: brute force over all triangles with the reference's acceptance rule
// (path_tracer.cu:83: dist < best && dist > Delta); exact-t ties resolved to the lowest primitive ID
// (SURVEY §8a tie-break rule) — which ascending iteration with strict '<' gives.
Hit brute_force(const Ray& ray, const float* wverts, const uint32_t* indices, int F) {
    Hit h = no_hit();
    for (int f = 0; f < F; ++f) {
        float dist, u, v;
        if (ray_hit_triangle(ray, ld3(wverts, indices[3 * f]), ld3(wverts, indices[3 * f + 1]), ld3(wverts, indices[3 * f + 2]),
                             dist, u, v)) {
            if (dist < h.dist && dist > kDelta) {
                h.dist = dist;
                h.idx = f;
                h.u = u;
                h.v = v;
            }
        }
    }
    return h;
}

// ----------------------------------------------------------------------------------------------
// RNG — the repo's counter-based generator (DESIGN.md "RNG spec"); replaces include/sampler.h:12-15
// (cuRAND XORWOW).  Uniforms are in (0,1] like curand_uniform.
// ----------------------------------------------------------------------------------------------
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline uint64_t path_key(uint64_t seed, uint32_t pixel, uint32_t sample) {
    return mix64(mix64(seed + kGolden) ^ (((uint64_t)pixel << 32) | (uint64_t)sample));
}
inline float u01(uint32_t x) { return (float)((x >> 8) + 1u) * (1.0f / 16777216.0f); }
inline void draw_pair(uint64_t key, uint32_t pairIdx, float& a, float& b) {
    uint64_t z = pairIdx == 0u ? key : mix64(key + (uint64_t)pairIdx * kGolden);
    a = u01((uint32_t)(z >> 32));
    b = u01((uint32_t)z);
}

// ----------------------------------------------------------------------------------------------
// src/path_tracer.cu — ray generation, BSDF sampling, lights, estimator
// ----------------------------------------------------------------------------------------------
struct Camera {
    float vfov, aspect, nearPlane;
    float c2w[16];
};

// path_tracer.cu:42-59 with the two uniforms passed in.
Ray sample_ray(float ju, float jv, const Camera& cam, int px, int py, int width, int height) {
    float plx = ju + (float)px, ply = jv + (float)py;   // Vec2 += Vec2(pixelX, pixelY)
    plx = plx * (1.0f / width);
    ply = ply * (1.0f / height);
    float tanHalf = tanf(cam.vfov * 0.5f);
    float sensorH = 2.0f * tanHalf;
    float sensorW = cam.aspect * sensorH;
    plx = plx * sensorW;
    ply = ply * sensorH;
    V3 dir = v3(plx, ply, 0.0f) - v3(0.5f * sensorW, 0.5f * sensorH, 1.0f);
    dir = normalize(xform_dir(cam.c2w, dir));
    V3 origin = xform_point(cam.c2w, v3(0.0f, 0.0f, 0.0f));
    return Ray{origin, dir};
}

inline int mtl_linear_search(int fid, const MtlInterval* lut, int objCnt) {   // path_tracer.cu:125-135
    int i = objCnt - 1;
    for (; i >= 0; --i)
        if (fid >= lut[i].begin) break;
    if (i < 0) i = 0;   // reference reads lut[-1] (UB) when fid < lut[0].begin; cannot happen (begin of object 0 is 0)
    return lut[i].mtl;
}
inline V3 reflect(V3 dir, V3 n) { return dir - 2.0f * dot(dir, n) * n; }   // path_tracer.cu:137-141
// path_tracer.cu:143-163
inline V3 refract(V3 dir, V3 n, float ior, float& cosI, float& eta, bool& tir) {
    cosI = dot(dir, n);
    eta = cosI > 0.0f ? ior : 1.0f / ior;
    n = cosI > 0.0f ? -n : n;
    cosI = fabsf(cosI);
    float sin2I = 1.0f - cosI * cosI;
    float sin2T = eta * eta * sin2I;
    if (sin2T >= 1.0f) {
        tir = true;
        return v3(0, 0, 0);
    }
    float cosT = sqrtf(1.0f - sin2T);
    return eta * dir + (cosI * eta - cosT) * n;
}
inline float schlick_fresnel(float cosI, float eta) {   // path_tracer.cu:165-173
    float f0 = (1.0f - eta) / (1.0f + eta);
    f0 *= f0;
    float m = clampf(1.0f - cosI, 1.0f, 0.0f);
    float m2 = m * m;
    return f0 + (1.0f - f0) * m2 * m2 * m;
}
// include/sampler.h:75-89 with (xi1 -> phi, xi2 -> cos^2) passed in draw order.
inline V3 hemisphere_cosine(float xi1, float xi2, V3 n) {
    V3 xBase = n.z == 0.0f ? v3(0, 0, 1) : v3(1.0f, 0.0f, -n.x / n.z);
    xBase = xBase / norm(xBase);
    V3 zBase = cross(xBase, n);
    float phi = 2.0f * kPi * xi1;
    float cosT = sqrtf(xi2);
    float sinT = sqrtf(1.0f - cosT * cosT);
    float x = cosf(phi) * sinT;
    float z = sinf(phi) * sinT;
    float y = cosT;
    return x * xBase + y * n + z * zBase;
}
inline float hemisphere_cosine_pdf(V3 dir, V3 n) {   // sampler.h:91-96
    float c = dot(dir, n);
    float f = c > 0.0f ? 1.0f : 0.0f;
    return (c / kPi) * f;
}
// NOT in the reference (TPT_RENDER_ROUGH_METALS): GGX reflection for metals with alpha = roughness^2 >= 1e-3,
// half vector from the distribution of visible normals (Heitz 2018), weight G1(wi) on top of baseColor; same local frame as
// hemisphere_cosine.  Returns G1(wi) (0 below the surface).
inline float rough_metal_direction(V3 rayDir, V3 n, float alpha, float xi1, float xi2, V3& out) {
    V3 nn = dot(rayDir, n) > 0.0f ? n * -1.0f : n;
    V3 xBase = nn.z == 0.0f ? v3(0, 0, 1) : v3(1.0f, 0.0f, -nn.x / nn.z);
    xBase = xBase / norm(xBase);
    V3 zBase = cross(xBase, nn);
    V3 wo = v3(-dot(rayDir, xBase), -dot(rayDir, zBase), -dot(rayDir, nn));
    V3 vh = v3(alpha * wo.x, alpha * wo.y, wo.z);
    vh = vh / norm(vh);
    float lensq = vh.x * vh.x + vh.y * vh.y;
    V3 t1 = lensq > 0.0f ? v3(-vh.y, vh.x, 0.0f) / sqrtf(lensq) : v3(1, 0, 0);
    V3 t2 = cross(vh, t1);
    float r = sqrtf(xi1), phi = 2.0f * kPi * xi2;
    float p1 = r * cosf(phi), sh = 0.5f * (1.0f + vh.z);
    float p2 = (1.0f - sh) * sqrtf(std::max(1.0f - p1 * p1, 0.0f)) + sh * (r * sinf(phi));
    V3 nh = p1 * t1 + p2 * t2 + sqrtf(std::max(1.0f - p1 * p1 - p2 * p2, 0.0f)) * vh;
    V3 ne = v3(alpha * nh.x, alpha * nh.y, std::max(nh.z, 0.0f));
    ne = ne / norm(ne);
    V3 wi = (2.0f * dot(wo, ne)) * ne - wo;
    out = wi.x * xBase + wi.y * zBase + wi.z * nn;
    if (!(wi.z > 0.0f)) return 0.0f;
    return 2.0f * wi.z / (wi.z + sqrtf(alpha * alpha + (1.0f - alpha * alpha) * wi.z * wi.z));
}
// path_tracer.cu:187-225.  xi1/xi2 are the pair of uniforms reserved for this call (a dielectric
// consumes xi1 only, a mirror none, a diffuse surface both, in that order).
inline float get_new_direction(V3 rayDir, V3 n, const Material& m, float xi1, float xi2, V3& nextDir, float& attenFactor,
                               bool roughMetals = false) {
    float prob;
    const float alpha = m.unused[2] * m.unused[2];   // roughness (include/material.h:113), read only by the rough-metal extension
    if (roughMetals && m.metallic > 0.0f && !(m.eta > 0.0f) && alpha >= 1e-3f) {
        attenFactor = rough_metal_direction(rayDir, n, alpha, xi1, xi2, nextDir);
        prob = 1.0f;
    } else if (m.eta > 0.0f) {
        bool tir = false;
        float eta, cosI;
        V3 refr = refract(rayDir, n, m.eta, cosI, eta, tir);
        V3 refl = reflect(rayDir, n);
        float fr = tir ? 1.0f : schlick_fresnel(cosI, eta);
        nextDir = (xi1 < fr) ? refl : refr;   // sampler.h:98-101 CoinFlip
        prob = 1.0f;
        attenFactor = 1.0f;
    } else if (m.metallic > 0.0f) {
        attenFactor = 1.0f;
        nextDir = reflect(rayDir, n);
        prob = 1.0f;
    } else {
        float sign = dot(rayDir, n) > 0.0f ? -1.0f : 1.0f;
        n = n * sign;
        nextDir = hemisphere_cosine(xi1, xi2, n);
        attenFactor = fabsf(dot(nextDir, n)) / kPi;
        prob = hemisphere_cosine_pdf(nextDir, n);
    }
    return prob;
}

struct Incoming {   // delta_light.h:16-22
    V3 radiance, direction;
    float distance;
};
inline void calc_dist_attenuation(Incoming& in) {   // delta_light.h:25-33
    const float invRadiusSqr = 0.01f;
    float d2 = in.distance * in.distance;
    float att = 1.0f / (d2 + 1.0f);
    att *= square(saturate(1.0f - square(d2 * invRadiusSqr)));
    in.radiance = in.radiance * att;
}
inline Incoming light_sample(const DeltaLight& L, V3 p) {   // delta_light.h:38-46,56-63,73-85,108-127
    Incoming r;
    r.radiance = v3(0, 0, 0);
    r.direction = v3(0, 0, 0);
    r.distance = 0.0f;
    V3 color = v3(L.color[0], L.color[1], L.color[2]);
    if (L.type == 0) {
        r.direction = v3(L.a[0], L.a[1], L.a[2]) - p;
        r.distance = norm(r.direction);
        r.direction = r.direction / r.distance;
        r.radiance = color * L.intensity;
    } else if (L.type == 1) {
        r.direction = -v3(L.a[0], L.a[1], L.a[2]);
        r.radiance = color * L.intensity;
        r.distance = 0.0f;
    } else if (L.type == 2) {
        r.direction = v3(L.a[0], L.a[1], L.a[2]) - p;
        r.distance = norm(r.direction);
        r.direction = r.direction / r.distance;
        float cosT = dot(-r.direction, v3(L.b[0], L.b[1], L.b[2]));
        float fall = square(saturate(cosT - L.cosOuter) * L.invConeDiff);
        r.radiance = color * L.intensity * fall;
    }
    calc_dist_attenuation(r);
    return r;
}

// include/env_light.cuh:72-78
inline void vec2uv(V3 d, float& u, float& v) {
    u = atan2f(d.z, d.x) / (2.0f * kPi);
    if (u < 0.0f) u += 1.0f;
    v = 1.0f - acosf(clampf(d.y, 1.0f, -1.0f)) / kPi;
}
// path_tracer.cu:288-294 + texture.cu:160-168: tex2DLod<uchar4>, normalised coordinates, point
// filter, clamp addressing, level 0: texel = clamp(floor(u*W)), clamp(floor(v*H)).  Row 0 of the
// texture is the bottom scanline of the file (FreeImage is bottom-up, include/picture.h:41-42).
// `rgbf` (not in the reference, whose Texture is 8-bit only): linear float RGB, 3 floats per texel, same lookup — the HDR
// maps of SURVEY §8f-2.
inline V3 sample_env(const uint8_t* rgba, int ew, int eh, V3 dir, const float* rgbf = nullptr) {
    if (!rgba && !rgbf) return v3(0, 0, 0);
    float u, v;
    vec2uv(dir, u, v);
    int x = (int)floorf(u * (float)ew), y = (int)floorf(v * (float)eh);
    x = std::min(std::max(x, 0), ew - 1);
    y = std::min(std::max(y, 0), eh - 1);
    if (rgbf) {
        const float* t = rgbf + 3 * ((size_t)y * ew + x);
        return v3(t[0], t[1], t[2]);
    }
    const uint8_t* t = rgba + 4 * ((size_t)y * ew + x);
    return v3((float)t[0], (float)t[1], (float)t[2]) / 255.0f;
}

struct SceneView {
    const uint32_t* indices;
    int F;
    const float* wverts;
    const float* wnormals;
    const Material* materials;
    int M;
    const MtlInterval* lut;
    int O;
    const DeltaLight* lights;
    int L;
    const RefNode* nodes;
    const uint8_t* env;
    int ew, eh;
    const float* envf = nullptr;   // float RGB map instead of `env` (HDR)
    bool roughMetals = false;      // TPT_RENDER_ROUGH_METALS
};

struct Counters {
    uint64_t rays, nInt, nLeaf;
};

constexpr int kMaxDepth = 64;

// path_tracer.cu:346-432 — one sample of the estimator, backward unwinding exactly as written.
V3 trace_sample(const SceneView& S, const Camera& cam, int px, int py, int W, int H, int maxDepth, uint64_t seed,
                uint32_t sampleIdx, Counters* C) {
    uint32_t pixel = (uint32_t)(px + py * W);
    uint64_t key = path_key(seed, pixel, sampleIdx);
    float ju, jv;
    draw_pair(key, 0, ju, jv);
    Ray ray = sample_ray(ju, jv, cam, px, py, W, H);

    float pStack[kMaxDepth];
    int mtlIdxStack[kMaxDepth];
    V3 directLightStack[kMaxDepth];
    V3 attenuation[kMaxDepth];

    auto traverse = [&](const Ray& r, Hit& h) {
        traverse_reference(r, S.nodes, S.wverts, S.indices, S.F, h, C ? &C->nInt : nullptr, C ? &C->nLeaf : nullptr);
        if (C) ++C->rays;
    };

    int depth = 0;
    V3 radiance = v3(0, 0, 0);
    for (; depth < maxDepth; depth += 1) {
        Hit status = no_hit();
        traverse(ray, status);
        if (status.idx < 0) {
            radiance = sample_env(S.env, S.ew, S.eh, ray.d, S.envf);
            break;
        }
        uint32_t i0 = S.indices[status.idx * 3 + 0], i1 = S.indices[status.idx * 3 + 1], i2 = S.indices[status.idx * 3 + 2];
        float u = status.u, v = status.v, w = 1.0f - u - v;
        V3 normal = normalize(w * ld3(S.wnormals, i0) + u * ld3(S.wnormals, i1) + v * ld3(S.wnormals, i2));
        ray.o = ray.o + status.dist * ray.d;
        int mtlIndex = mtl_linear_search(status.idx, S.lut, S.O);
        const Material& mtl = S.materials[mtlIndex];
        V3 base = v3(mtl.base[0], mtl.base[1], mtl.base[2]);

        float xi1, xi2;
        draw_pair(key, 1u + 2u * (uint32_t)depth, xi1, xi2);
        V3 newDir;
        float attenFactor;
        float prob = get_new_direction(ray.d, normal, mtl, xi1, xi2, newDir, attenFactor, S.roughMetals);
        attenuation[depth] = base * attenFactor;
        pStack[depth] = prob;

        // sampleDeltaLights, path_tracer.cu:265-286
        V3 directRadiance = v3(0, 0, 0);
        for (int i = 0; i < S.L; ++i) {
            Incoming in = light_sample(S.lights[i], ray.o);
            Ray shadow{ray.o, in.direction};
            Hit sh = no_hit();
            traverse(shadow, sh);
            if (sh.idx == -1) directRadiance = directRadiance + base * in.radiance;
        }
        if (!(mtl.eta >= 1.0f || mtl.metallic > 0.0f)) {
            float d1, d2;
            draw_pair(key, 2u + 2u * (uint32_t)depth, d1, d2);
            V3 directDir;
            float af;
            get_new_direction(ray.d, normal, mtl, d1, d2, directDir, af);
            Ray dray{ray.o, directDir};
            Hit dh = no_hit();
            traverse(dray, dh);
            ray.d = directDir;
            if (dh.idx >= 0) {
                int dm = mtl_linear_search(dh.idx, S.lut, S.O);
                float e = S.materials[dm].emission;
                directLightStack[depth] = v3(1.0f, 1.0f, 1.0f) * v3(e, e, e) + directRadiance;
            } else {
                directLightStack[depth] = directRadiance;
            }
        } else {
            directLightStack[depth] = directRadiance;
        }
        mtlIdxStack[depth] = mtlIndex;
        if (mtl.emission > 0.0f) {
            depth++;
            break;
        }
        ray.d = newDir;
    }
    depth -= 1;
    while (depth >= 0) {
        const Material& mtl = S.materials[mtlIdxStack[depth]];
        if (mtl.emission > 0.0f) {
            radiance = v3(1.0f, 1.0f, 1.0f) * mtl.emission;
        } else {
            radiance = ((directLightStack[depth] + radiance) * attenuation[depth]) / pStack[depth];
        }
        depth -= 1;
    }
    return radiance;
}

}   // namespace

// ================================================================================================
// C entry points (ctypes)
// ================================================================================================
extern "C" {

float orc_frsqrt(float x) { return frsqrt(x); }
float orc_cosf(float x) { return cosf(x); }   // std::cos(float) as used at src/mesh.cu:294-295

int64_t orc_morton(float x, float y, float z) { return morton_code(v3(x, y, z)); }

int orc_ray_hit_bbox(const float* ray6, const float* mn, const float* mx) {
    Ray r{v3(ray6[0], ray6[1], ray6[2]), v3(ray6[3], ray6[4], ray6[5])};
    return ray_hit_bbox(r, mn, mx) ? 1 : 0;
}

int orc_ray_hit_triangle(const float* ray6, const float* tri9, float* out3) {
    Ray r{v3(ray6[0], ray6[1], ray6[2]), v3(ray6[3], ray6[4], ray6[5])};
    float d = 0, u = 0, v = 0;
    bool ok = ray_hit_triangle(r, ld3(tri9, 0), ld3(tri9, 1), ld3(tri9, 2), d, u, v);
    out3[0] = d;
    out3[1] = u;
    out3[2] = v;
    return ok ? 1 : 0;
}

void orc_reflect(const float* d, const float* n, float* out) {
    V3 r = reflect(ld3(d, 0), ld3(n, 0));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// out: dir[3], cosI, eta, tir
void orc_refract(const float* d, const float* n, float ior, float* out6) {
    float cosI, eta;
    bool tir = false;
    V3 r = refract(ld3(d, 0), ld3(n, 0), ior, cosI, eta, tir);
    out6[0] = r.x; out6[1] = r.y; out6[2] = r.z;
    out6[3] = cosI; out6[4] = eta; out6[5] = tir ? 1.0f : 0.0f;
}
float orc_schlick(float cosI, float eta) { return schlick_fresnel(cosI, eta); }

void orc_hemisphere_cosine(float xi1, float xi2, const float* n, float* out) {
    V3 r = hemisphere_cosine(xi1, xi2, ld3(n, 0));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

// out: radiance[3], direction[3], distance
void orc_light_sample(const void* light52, const float* p, float* out7) {
    Incoming in = light_sample(*reinterpret_cast<const DeltaLight*>(light52), ld3(p, 0));
    out7[0] = in.radiance.x; out7[1] = in.radiance.y; out7[2] = in.radiance.z;
    out7[3] = in.direction.x; out7[4] = in.direction.y; out7[5] = in.direction.z;
    out7[6] = in.distance;
}

void orc_vec2uv(const float* d, float* uv) { vec2uv(ld3(d, 0), uv[0], uv[1]); }

void orc_sample_env(const uint8_t* rgba, int ew, int eh, const float* d, float* rgb) {
    V3 r = sample_env(rgba, ew, eh, ld3(d, 0));
    rgb[0] = r.x; rgb[1] = r.y; rgb[2] = r.z;
}

void orc_rng_pair(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t pairIdx, float* out2) {
    draw_pair(path_key(seed, pixel, sample), pairIdx, out2[0], out2[1]);
}

// include/transform.h:28-33 localToParent = Translate(loc) * Mat4(RotateFromQuat(q)) * Scale(s);
// include/math/quat.h:52-69; include/math/mat.h:169-176,282-298.  q = (x,y,z,w).
void orc_trs_to_mat4(const float* t, const float* q, const float* s, float* out16) {
    float x = q[0], y = q[1], z = q[2], w = q[3];
    float x2 = x * x, y2 = y * y, z2 = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    float R[16] = {1.0f - 2.0f * (y2 + z2), 2.0f * (xy + wz), 2.0f * (xz - wy), 0.0f,
                   2.0f * (xy - wz), 1.0f - 2.0f * (x2 + z2), 2.0f * (yz + wx), 0.0f,
                   2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (x2 + y2), 0.0f,
                   0.0f, 0.0f, 0.0f, 1.0f};
    float T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, t[0], t[1], t[2], 1.0f};
    float S[16] = {s[0], 0, 0, 0, 0, s[1], 0, 0, 0, 0, s[2], 0, 0, 0, 0, 1.0f};
    float TR[16];
    mat4_mul(T, R, TR);
    mat4_mul(TR, S, out16);
}

// src/mesh.cu:371-378 normal_to_world = (upper-left 3x3 as Mat4).transpose().inverse() with the
// cofactor inverse and 24-term determinant of include/math/mat.h:203-281.
void orc_normal_matrix(const float* l2w, float* out16) {
    float m[16] = {l2w[0], l2w[1], l2w[2], 0.0f, l2w[4], l2w[5], l2w[6], 0.0f, l2w[8], l2w[9], l2w[10], 0.0f, 0.0f, 0.0f, 0.0f, 1.0f};
    float c[4][4];   // c[col][row] of the transpose
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) c[j][i] = m[i * 4 + j];
    float r[4][4];
    r[0][0] = c[1][2] * c[2][3] * c[3][1] - c[1][3] * c[2][2] * c[3][1] + c[1][3] * c[2][1] * c[3][2] - c[1][1] * c[2][3] * c[3][2] - c[1][2] * c[2][1] * c[3][3] + c[1][1] * c[2][2] * c[3][3];
    r[0][1] = c[0][3] * c[2][2] * c[3][1] - c[0][2] * c[2][3] * c[3][1] - c[0][3] * c[2][1] * c[3][2] + c[0][1] * c[2][3] * c[3][2] + c[0][2] * c[2][1] * c[3][3] - c[0][1] * c[2][2] * c[3][3];
    r[0][2] = c[0][2] * c[1][3] * c[3][1] - c[0][3] * c[1][2] * c[3][1] + c[0][3] * c[1][1] * c[3][2] - c[0][1] * c[1][3] * c[3][2] - c[0][2] * c[1][1] * c[3][3] + c[0][1] * c[1][2] * c[3][3];
    r[0][3] = c[0][3] * c[1][2] * c[2][1] - c[0][2] * c[1][3] * c[2][1] - c[0][3] * c[1][1] * c[2][2] + c[0][1] * c[1][3] * c[2][2] + c[0][2] * c[1][1] * c[2][3] - c[0][1] * c[1][2] * c[2][3];
    r[1][0] = c[1][3] * c[2][2] * c[3][0] - c[1][2] * c[2][3] * c[3][0] - c[1][3] * c[2][0] * c[3][2] + c[1][0] * c[2][3] * c[3][2] + c[1][2] * c[2][0] * c[3][3] - c[1][0] * c[2][2] * c[3][3];
    r[1][1] = c[0][2] * c[2][3] * c[3][0] - c[0][3] * c[2][2] * c[3][0] + c[0][3] * c[2][0] * c[3][2] - c[0][0] * c[2][3] * c[3][2] - c[0][2] * c[2][0] * c[3][3] + c[0][0] * c[2][2] * c[3][3];
    r[1][2] = c[0][3] * c[1][2] * c[3][0] - c[0][2] * c[1][3] * c[3][0] - c[0][3] * c[1][0] * c[3][2] + c[0][0] * c[1][3] * c[3][2] + c[0][2] * c[1][0] * c[3][3] - c[0][0] * c[1][2] * c[3][3];
    r[1][3] = c[0][2] * c[1][3] * c[2][0] - c[0][3] * c[1][2] * c[2][0] + c[0][3] * c[1][0] * c[2][2] - c[0][0] * c[1][3] * c[2][2] - c[0][2] * c[1][0] * c[2][3] + c[0][0] * c[1][2] * c[2][3];
    r[2][0] = c[1][1] * c[2][3] * c[3][0] - c[1][3] * c[2][1] * c[3][0] + c[1][3] * c[2][0] * c[3][1] - c[1][0] * c[2][3] * c[3][1] - c[1][1] * c[2][0] * c[3][3] + c[1][0] * c[2][1] * c[3][3];
    r[2][1] = c[0][3] * c[2][1] * c[3][0] - c[0][1] * c[2][3] * c[3][0] - c[0][3] * c[2][0] * c[3][1] + c[0][0] * c[2][3] * c[3][1] + c[0][1] * c[2][0] * c[3][3] - c[0][0] * c[2][1] * c[3][3];
    r[2][2] = c[0][1] * c[1][3] * c[3][0] - c[0][3] * c[1][1] * c[3][0] + c[0][3] * c[1][0] * c[3][1] - c[0][0] * c[1][3] * c[3][1] - c[0][1] * c[1][0] * c[3][3] + c[0][0] * c[1][1] * c[3][3];
    r[2][3] = c[0][3] * c[1][1] * c[2][0] - c[0][1] * c[1][3] * c[2][0] - c[0][3] * c[1][0] * c[2][1] + c[0][0] * c[1][3] * c[2][1] + c[0][1] * c[1][0] * c[2][3] - c[0][0] * c[1][1] * c[2][3];
    r[3][0] = c[1][2] * c[2][1] * c[3][0] - c[1][1] * c[2][2] * c[3][0] - c[1][2] * c[2][0] * c[3][1] + c[1][0] * c[2][2] * c[3][1] + c[1][1] * c[2][0] * c[3][2] - c[1][0] * c[2][1] * c[3][2];
    r[3][1] = c[0][1] * c[2][2] * c[3][0] - c[0][2] * c[2][1] * c[3][0] + c[0][2] * c[2][0] * c[3][1] - c[0][0] * c[2][2] * c[3][1] - c[0][1] * c[2][0] * c[3][2] + c[0][0] * c[2][1] * c[3][2];
    r[3][2] = c[0][2] * c[1][1] * c[3][0] - c[0][1] * c[1][2] * c[3][0] - c[0][2] * c[1][0] * c[3][1] + c[0][0] * c[1][2] * c[3][1] + c[0][1] * c[1][0] * c[3][2] - c[0][0] * c[1][1] * c[3][2];
    r[3][3] = c[0][1] * c[1][2] * c[2][0] - c[0][2] * c[1][1] * c[2][0] + c[0][2] * c[1][0] * c[2][1] - c[0][0] * c[1][2] * c[2][1] - c[0][1] * c[1][0] * c[2][2] + c[0][0] * c[1][1] * c[2][2];
    float det = c[0][3] * c[1][2] * c[2][1] * c[3][0] - c[0][2] * c[1][3] * c[2][1] * c[3][0] - c[0][3] * c[1][1] * c[2][2] * c[3][0] + c[0][1] * c[1][3] * c[2][2] * c[3][0] +
                c[0][2] * c[1][1] * c[2][3] * c[3][0] - c[0][1] * c[1][2] * c[2][3] * c[3][0] - c[0][3] * c[1][2] * c[2][0] * c[3][1] + c[0][2] * c[1][3] * c[2][0] * c[3][1] +
                c[0][3] * c[1][0] * c[2][2] * c[3][1] - c[0][0] * c[1][3] * c[2][2] * c[3][1] - c[0][2] * c[1][0] * c[2][3] * c[3][1] + c[0][0] * c[1][2] * c[2][3] * c[3][1] +
                c[0][3] * c[1][1] * c[2][0] * c[3][2] - c[0][1] * c[1][3] * c[2][0] * c[3][2] - c[0][3] * c[1][0] * c[2][1] * c[3][2] + c[0][0] * c[1][3] * c[2][1] * c[3][2] +
                c[0][1] * c[1][0] * c[2][3] * c[3][2] - c[0][0] * c[1][1] * c[2][3] * c[3][2] - c[0][2] * c[1][1] * c[2][0] * c[3][3] + c[0][1] * c[1][2] * c[2][0] * c[3][3] +
                c[0][2] * c[1][0] * c[2][1] * c[3][3] - c[0][0] * c[1][2] * c[2][1] * c[3][3] - c[0][1] * c[1][0] * c[2][2] * c[3][3] + c[0][0] * c[1][1] * c[2][2] * c[3][3];
    float inv = 1.0f / det;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) out16[i * 4 + j] = r[i][j] * inv;
}

void orc_mat4_mul_point(const float* m, const float* p, float w, float* out3) {
    float v[4] = {p[0], p[1], p[2], w}, o[4];
    mat4_mul_vec4(m, v, o);
    out3[0] = o[0]; out3[1] = o[1]; out3[2] = o[2];
}

// src/path_tracer.cu:239-263 — object -> world vertices and (frsqrt-normalised) normals.  The
// reference does this per face corner with redundant identical writes; per vertex of each object
// gives the same values.  `lut[o].begin` is the first FACE of object o.
void orc_transform(const uint32_t* indices, int F, const float* verts, const float* normals, int V, const void* lut_, int O,
                   const float* vertTrans, const float* normalTrans, float* wverts, float* wnormals) {
    const MtlInterval* lut = reinterpret_cast<const MtlInterval*>(lut_);
    std::memset(wverts, 0, sizeof(float) * 3 * (size_t)V);
    std::memset(wnormals, 0, sizeof(float) * 3 * (size_t)V);
    for (int f = 0; f < F; ++f) {
        int o = O - 1;   // findIdxOfTrans, path_tracer.cu:227-237
        for (; o >= 0; --o)
            if (f >= lut[o].begin) break;
        if (o < 0) o = 0;
        for (int k = 0; k < 3; ++k) {
            uint32_t vid = indices[3 * f + k];
            V3 wv = xform_point(vertTrans + 16 * o, ld3(verts, vid));
            V3 wn = normalize(xform_dir(normalTrans + 16 * o, ld3(normals, vid)));
            wverts[3 * vid] = wv.x; wverts[3 * vid + 1] = wv.y; wverts[3 * vid + 2] = wv.z;
            wnormals[3 * vid] = wn.x; wnormals[3 * vid + 1] = wn.y; wnormals[3 * vid + 2] = wn.z;
        }
    }
}

int orc_build_reference_bvh(const float* wverts, const uint32_t* indices, int F, void* nodes36, int64_t* keys) {
    return build_reference_bvh(wverts, indices, F, reinterpret_cast<RefNode*>(nodes36), keys);
}

// cam: vfov, aspect, near, c2w[16] (19 floats).  useJitter == 0 -> pixel centres (0.5, 0.5);
// otherwise RNG pair 0 of (seed, pixel, sample 0).
void orc_primary_rays(const float* cam19, int W, int H, int useJitter, uint64_t seed, float* rays6) {
    Camera cam;
    cam.vfov = cam19[0]; cam.aspect = cam19[1]; cam.nearPlane = cam19[2];
    std::memcpy(cam.c2w, cam19 + 3, 64);
#pragma omp parallel for schedule(static)
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            float ju = 0.5f, jv = 0.5f;
            if (useJitter) draw_pair(path_key(seed, (uint32_t)(px + py * W), 0), 0, ju, jv);
            Ray r = sample_ray(ju, jv, cam, px, py, W, H);
            float* o = rays6 + 6 * ((size_t)py * W + px);
            o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
        }
}

// Brute force for LARGE triangle counts (the 10 M-triangle soup): the same Möller–Trumbore arithmetic and the same acceptance
// rule as brute_force() above, evaluated for 8 triangles at a time with the compiler's vector types.  Every vector operation is
// the element-wise IEEE operation of the scalar code (no contraction: -ffp-contract=off; no reassociation), so each lane
// computes bit for bit what ray_hit_triangle computes; lanes are then visited in ascending triangle order with the same strict
// `<`, which keeps the lowest-primitive-id tie rule.  Triangles are processed in cache-sized blocks (outer loop, ascending),
// all rays per block (parallel), so the cost is arithmetic, not memory.  tests/test_oracle_known_answers.py checks it equals
// mode 0 exactly.
typedef float vf8 __attribute__((vector_size(32)));
typedef int vi8 __attribute__((vector_size(32)));
constexpr int kBruteBlock = 2048;   // triangles per block: 9 SoA arrays x 2048 x 4 B = 72 KB

__attribute__((target_clones("avx2", "default")))
void brute_block(const float* soa, int fBase, int count, const float* rays6, long long n, Hit* out) {
    const int padded = (count + 7) & ~7;
    const float* V0x = soa, *V0y = soa + kBruteBlock, *V0z = soa + 2 * kBruteBlock;
    const float* E1x = soa + 3 * kBruteBlock, *E1y = soa + 4 * kBruteBlock, *E1z = soa + 5 * kBruteBlock;
    const float* E2x = soa + 6 * kBruteBlock, *E2y = soa + 7 * kBruteBlock, *E2z = soa + 8 * kBruteBlock;
#pragma omp parallel for schedule(dynamic, 64)
    for (long long i = 0; i < n; ++i) {
        const float ox = rays6[6 * i], oy = rays6[6 * i + 1], oz = rays6[6 * i + 2];
        const float dx = rays6[6 * i + 3], dy = rays6[6 * i + 4], dz = rays6[6 * i + 5];
        Hit h = out[i];
        for (int j = 0; j < padded; j += 8) {
            const vf8 v0x = *(const vf8*)(V0x + j), v0y = *(const vf8*)(V0y + j), v0z = *(const vf8*)(V0z + j);
            const vf8 e1x = *(const vf8*)(E1x + j), e1y = *(const vf8*)(E1y + j), e1z = *(const vf8*)(E1z + j);
            const vf8 e2x = *(const vf8*)(E2x + j), e2y = *(const vf8*)(E2y + j), e2z = *(const vf8*)(E2z + j);
            const vf8 tx = ox - v0x, ty = oy - v0y, tz = oz - v0z;                               // t = r.o - v0
            const vf8 px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;   // p = cross(r.d, e2)
            const vf8 qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;   // q = cross(t, e1)
            const vf8 denom = px * e1x + py * e1y + pz * e1z;
            const vf8 inv = 1.0f / denom;
            const vf8 u = (px * tx + py * ty + pz * tz) * inv;
            const vf8 v = (qx * dx + qy * dy + qz * dz) * inv;
            const vf8 dist = (qx * e2x + qy * e2y + qz * e2z) * inv;
            const vi8 reject = (denom == 0.0f) | (u < 0.0f) | (v < 0.0f) | ((u + v) > 1.0f);
            const vi8 ok = ~reject & (dist < h.dist) & (dist > kDelta);
            typedef long long vl4 __attribute__((vector_size(32)));
            const vl4 ok4 = (vl4)ok;
            if (((ok4[0] | ok4[1]) | (ok4[2] | ok4[3])) == 0) continue;
            for (int l = 0; l < 8; ++l)   // ascending ids, strict '<' against the running best: brute_force()'s rule
                if (ok[l] && j + l < count && dist[l] < h.dist) {
                    h.dist = dist[l];
                    h.idx = fBase + j + l;
                    h.u = u[l];
                    h.v = v[l];
                }
        }
        out[i] = h;
    }
}

void brute_force_blocked(const float* wverts, const uint32_t* indices, int F, const float* rays6, long long n, Hit* out) {
    for (long long i = 0; i < n; ++i) out[i] = no_hit();
    float* soa = static_cast<float*>(aligned_alloc(64, sizeof(float) * 9 * kBruteBlock));
    for (int f0 = 0; f0 < F; f0 += kBruteBlock) {
        const int count = std::min(kBruteBlock, F - f0);
        for (int k = 0; k < ((count + 7) & ~7); ++k) {
            V3 v0 = v3(0.f, 0.f, 0.f), e1 = v0, e2 = v0;   // padding lanes: denom == 0 -> rejected
            if (k < count) {
                const int f = f0 + k;
                v0 = ld3(wverts, indices[3 * f]);
                e1 = ld3(wverts, indices[3 * f + 1]) - v0;   // the subtractions ray_hit_triangle performs per test
                e2 = ld3(wverts, indices[3 * f + 2]) - v0;
            }
            soa[k] = v0.x; soa[kBruteBlock + k] = v0.y; soa[2 * kBruteBlock + k] = v0.z;
            soa[3 * kBruteBlock + k] = e1.x; soa[4 * kBruteBlock + k] = e1.y; soa[5 * kBruteBlock + k] = e1.z;
            soa[6 * kBruteBlock + k] = e2.x; soa[7 * kBruteBlock + k] = e2.y; soa[8 * kBruteBlock + k] = e2.z;
        }
        brute_block(soa, f0, count, rays6, n, out);
    }
    free(soa);
}

// mode 0: brute force (authority; lowest ID on exact ties).  mode 1: reference BVH traversal order.  mode 2: mode 0's
// result computed block-wise with vector arithmetic (brute_force_blocked; for millions of triangles).
// hits: (idx:int32, dist, u, v) 16 bytes each.  counters (optional): [rays, nInt, nLeaf].
int orc_trace_rays(const float* wverts, const uint32_t* indices, int F, const void* nodes36, const float* rays6, size_t n, int mode,
                   void* hits16, uint64_t* counters) {
    const RefNode* nodes = reinterpret_cast<const RefNode*>(nodes36);
    Hit* out = reinterpret_cast<Hit*>(hits16);
    int bad = 0;
    uint64_t nI = 0, nL = 0;
    if (mode == 2) {
        brute_force_blocked(wverts, indices, F, rays6, (long long)n, out);
        if (counters) counters[0] = n, counters[1] = 0, counters[2] = 0;
        return 0;
    }
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : bad, nI, nL)
    for (long long i = 0; i < (long long)n; ++i) {
        Ray r{ld3(rays6, 2 * i), ld3(rays6, 2 * i + 1)};
        if (mode == 0) {
            out[i] = brute_force(r, wverts, indices, F);
        } else {
            Hit h = no_hit();
            uint64_t a = 0, b = 0;
            if (!traverse_reference(r, nodes, wverts, indices, F, h, &a, &b)) bad += 1;
            nI += a;
            nL += b;
            out[i] = h;
        }
    }
    if (counters) {
        counters[0] = n;
        counters[1] = nI;
        counters[2] = nL;
    }
    return bad;
}

// Re-test one triangle with edge tolerance eps (classification aid for the parity harness: is a
// mismatching primitive a legitimate near-tie / edge-grazing hit?).  Returns 1 and t if the ray
// hits triangle `fid` with u,v >= -eps, u+v <= 1+eps, t > Delta.
int orc_retest_triangle(const float* wverts, const uint32_t* indices, int fid, const float* ray6, float eps, float* t_out) {
    Ray r{ld3(ray6, 0), ld3(ray6, 1)};
    V3 v0 = ld3(wverts, indices[3 * fid]), v1 = ld3(wverts, indices[3 * fid + 1]), v2 = ld3(wverts, indices[3 * fid + 2]);
    V3 e1 = v1 - v0, e2 = v2 - v0, t = r.o - v0, p = cross(r.d, e2), q = cross(t, e1);
    float denom = dot(p, e1);
    if (denom == 0.0f) return 0;
    float inv = 1.0f / denom, u = dot(p, t) * inv, v = dot(q, r.d) * inv, d = dot(q, e2) * inv;
    *t_out = d;
    return (u >= -eps && v >= -eps && u + v <= 1.0f + eps && d > kDelta * 0.5f) ? 1 : 0;
}

// src/path_tracer.cu:296-435 `trace` + doTrace's zeroed radiance (:534): out[pixel*3..] = sum over
// samples [sampleOffset, sampleOffset + spp) of the per-sample radiance, float accumulation in sample
// order exactly as the kernel does (totalRad += radiance; color += totalRad).
// Region [x0,x1) x [y0,y1) only (bounded CPU samples); pixels outside are left untouched.
// counters (optional): [rays, nInt, nLeaf].
static bool g_rough_metals = false;
// estimator extensions beyond the reference, off by default (bit 0: TPT_RENDER_ROUGH_METALS); applies to later orc_render* calls
void orc_set_extensions(int flags) { g_rough_metals = (flags & 1) != 0; }
void orc_render_env(const uint32_t* indices, int F, const float* wverts, const float* wnormals, const void* materials, int M,
                    const void* lut, int O, const void* lights, int L, const void* nodes36, const uint8_t* env, const float* envf,
                    int ew, int eh, const float* cam19, int W, int H, int x0, int y0, int x1, int y1, int spp, int sampleOffset,
                    int sampleStride, int maxDepth, uint64_t seed, float* out, uint64_t* counters);
void orc_render(const uint32_t* indices, int F, const float* wverts, const float* wnormals, const void* materials, int M,
                const void* lut, int O, const void* lights, int L, const void* nodes36, const uint8_t* env, int ew, int eh,
                const float* cam19, int W, int H, int x0, int y0, int x1, int y1, int spp, int sampleOffset, int sampleStride,
                int maxDepth, uint64_t seed, float* out, uint64_t* counters) {
    orc_render_env(indices, F, wverts, wnormals, materials, M, lut, O, lights, L, nodes36, env, nullptr, ew, eh, cam19, W, H, x0, y0,
                   x1, y1, spp, sampleOffset, sampleStride, maxDepth, seed, out, counters);
}
// the same with either texel format of the environment: `env` RGBA8 (the reference's) or `envf` float RGB (HDR)
void orc_render_env(const uint32_t* indices, int F, const float* wverts, const float* wnormals, const void* materials, int M,
                    const void* lut, int O, const void* lights, int L, const void* nodes36, const uint8_t* env, const float* envf,
                    int ew, int eh, const float* cam19, int W, int H, int x0, int y0, int x1, int y1, int spp, int sampleOffset,
                    int sampleStride, int maxDepth, uint64_t seed, float* out, uint64_t* counters) {
    SceneView S{indices, F, wverts, wnormals, reinterpret_cast<const Material*>(materials), M,
                reinterpret_cast<const MtlInterval*>(lut), O, reinterpret_cast<const DeltaLight*>(lights), L,
                reinterpret_cast<const RefNode*>(nodes36), env, ew, eh, envf, g_rough_metals};
    Camera cam;
    cam.vfov = cam19[0]; cam.aspect = cam19[1]; cam.nearPlane = cam19[2];
    std::memcpy(cam.c2w, cam19 + 3, 64);
    if (maxDepth > kMaxDepth) maxDepth = kMaxDepth;
    uint64_t cr = 0, ci = 0, cl = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : cr, ci, cl)
    for (int py = y0; py < y1; ++py) {
        Counters C{0, 0, 0};
        for (int px = x0; px < x1; ++px) {
            V3 total = v3(0, 0, 0);
            for (int s = 0; s < spp; ++s) {
                V3 r = trace_sample(S, cam, px, py, W, H, maxDepth, seed, (uint32_t)(sampleOffset + s * sampleStride), counters ? &C : nullptr);
                total = total + r;
            }
            float* o = out + 3 * ((size_t)py * W + px);
            o[0] += total.x; o[1] += total.y; o[2] += total.z;
        }
        cr += C.rays; ci += C.nInt; cl += C.nLeaf;
    }
    if (counters) {
        counters[0] = cr; counters[1] = ci; counters[2] = cl;
    }
}

// src/path_tracer.cu:451-471 copyToFB + include/material.h:74-81 toUChar: BGR(A untouched), row
// H-1-y, truncating conversion of clamp(255 * L/spp).
void orc_resolve_bgra8(const float* radiance, int W, int H, int spp, uint8_t* fb) {
    float inv = 1.0f / (float)spp;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const float* c = radiance + 3 * ((size_t)y * W + x);
            uint8_t* p = fb + 4 * ((size_t)(H - y - 1) * W + x);
            float r = c[0] * inv, g = c[1] * inv, b = c[2] * inv;
            p[0] = (uint8_t)clampf(b * 255.0f, 255.0f, 0.0f);
            p[1] = (uint8_t)clampf(g * 255.0f, 255.0f, 0.0f);
            p[2] = (uint8_t)clampf(r * 255.0f, 255.0f, 0.0f);
        }
}

int orc_num_threads() {
#if defined(_OPENMP)
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}   // extern "C"
