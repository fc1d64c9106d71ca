// traverse.cuh — software BVH traversal (B200 has no RT cores).
//
// Replaces traverseBVH / rayHitBBox / rayHitTriangle of the reference (src/path_tracer.cu:61-107,
// include/geometry_queries.h:18-46,66-86).  Same acceptance rule for a hit — Möller–Trumbore,
// two-sided, u >= 0, v >= 0, u+v <= 1, t > 2e-4, strictly closer than the best so far — but:
//   * while-while loop structure; ordered (near child first) with the far child pushed, and boxes are culled against the current
//     best t (the reference visits every box the infinite LINE touches);
//   * one reciprocal per ray, not three divisions per box;
//   * 64-B nodes read with four 128-bit loads, both child boxes in the parent;
//   * exact-t ties resolve to the LOWEST primitive id, so the result does not depend on tree shape
//     (the reference keeps the first triangle met in its own traversal order).
#pragma once
#include "scene.cuh"

namespace tpt {

constexpr int kStackSize = 96;   // LBVH depth <= 63 key bits + 32 index tie-break bits

struct HitRec {
    float t, u, v;
    int leaf;    // sorted triangle index, -1 = miss
    int prim;    // original face id
};

__device__ __forceinline__ float safe_rcp(float d) {
    // A zero direction component makes the slab parallel: substitute a tiny value of the same sign
    // so that (b - o) * inv is +-huge with the right sign (or 0 when the origin lies on the plane).
    // MUFU.RCP (<= 1 ulp) instead of an IEEE division: the reciprocal only feeds the conservative box
    // test, whose 2^-21 widening covers sub + mul + rcp rounding (3 * 2^-24 + 2^-23 < 2^-21).
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fabsf(d) > 1e-30f ? d : copysignf(1e-30f, d)));
    return r;
}

// 256-bit read-only load (sm_100: LDG.E.ENL2.256).  A 64-B node is two of these instead of three 128-bit
// loads and a 64-bit one: for incoherent rays every load instruction costs one L1 wavefront per distinct
// line a lane touches, so halving the instruction count halves the L1 data-pipe time of a box step.
struct F8 {
    float4 a, b;
};
__device__ __forceinline__ F8 ldg256(const float4* p) {
    F8 r;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
        : "l"(p));
    return r;
}

// Conservative slab test for one child: returns entry distance in `lo`.  The exit distance is
// widened and the entry distance narrowed by 2^-21 relative so that rounding in (b-o)*inv can
// never cull a box whose triangle the Möller–Trumbore test would accept.
__device__ __forceinline__ bool slab(float lox, float hix, float loy, float hiy, float loz, float hiz, float3 o, float3 inv,
                                     float tbest, float& lo) {
    float t0x = (lox - o.x) * inv.x, t1x = (hix - o.x) * inv.x;
    float t0y = (loy - o.y) * inv.y, t1y = (hiy - o.y) * inv.y;
    float t0z = (loz - o.z) * inv.z, t1z = (hiz - o.z) * inv.z;
    float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    tn = fmaxf(tn * 0.9999995f, 0.0f);
    tf = fminf(tf * 1.0000005f, tbest);
    lo = tn;
    return tn <= tf;
}

// One Möller–Trumbore test exactly as geometry_queries.h:66-86 evaluates it (one rounding per operation),
// followed by the acceptance rule of src/path_tracer.cu:83 with the lowest-primitive-id tie-break.
__device__ __forceinline__ bool leaf_test(const BvhView& bvh, int leaf, float3 o, float3 d, HitRec& best) {
    const float4* tp = bvh.tri_geom + 3 * (size_t)leaf;
    const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
    const float3 v0 = f3(a.x, a.y, a.z), e1 = f3(a.w, b.x, b.y), e2 = f3(b.z, b.w, c.x);
    const float3 tv = xsub3(o, v0);
    const float3 p = xcross(d, e2);
    const float3 q = xcross(tv, e1);
    const float denom = xdot(p, e1);
    if (denom == 0.0f) return false;
    const float invDenom = __fdiv_rn(1.0f, denom);
    const float u = xmul(xdot(p, tv), invDenom);
    const float v = xmul(xdot(q, d), invDenom);
    if (u < 0.0f || v < 0.0f || xadd(u, v) > 1.0f) return false;
    const float t = xmul(xdot(q, e2), invDenom);
    const int prim = __float_as_int(c.y);
    if (t > kDelta && (t < best.t || (t == best.t && prim < best.prim))) {
        best.t = t;
        best.u = u;
        best.v = v;
        best.leaf = leaf;
        best.prim = prim;
        return true;
    }
    return false;
}

constexpr int kStackSentinel = (int)0x80000000;   // negative, and never a valid leaf code (~leaf >= -2^30)
// The newest stack entry can live in a REGISTER (`top`); memory then only sees an entry when a second one is pushed on top of it.  In a
// depth-first walk most pushes are popped again before the next push (visit: push far child, descend near child, miss, pop), so
// about half of the stack's loads and stores — each a wavefront of the L1 data pipe per lane — never happen.  Measured (ms per frame
// stage / Mrays/s, with | without): the while-while walk of k_primary gains (ball 16.59 | 16.92, tir 2.42 | 2.59), the per-lane state
// machine of k_wave and k_trace_rays_persistent LOSES (box 9.60 | 9.08, ball 11.94 | 11.85, soup 1 240 | 1 277 Mrays/s: at 64 / 48
// registers the extra live value spills).  So: on in traverse(), off in the state machine.  kNoTop: the register is empty.
#ifndef TPT_STACK_TOP
#define TPT_STACK_TOP 1      // traverse() (while-while)
#endif
#ifndef TPT_STACK_TOP_SM
#define TPT_STACK_TOP_SM 0   // TravState (persistent kernels)
#endif
constexpr int kNoTop = 0x7fffffff;

// "while-while" traversal (Aila & Laine 2009): the inner loop descends internal nodes until the lane holds
// a leaf, so that the lanes of a warp run their (expensive, exactly-rounded) triangle tests together
// instead of one lane's leaf serialising against the other lanes' box tests.
// `stack` is caller-owned (kStackSize ints) with stack[0] == kStackSentinel already written: a kernel that
// traces many rays per thread initialises it once.
template <bool COUNT>
__device__ __forceinline__ HitRec traverse(const BvhView& bvh, float3 o, float3 d, bool anyHit, int* stack, unsigned long long* nodeVisits,
                                           unsigned long long* triTests) {
    HitRec best;
    best.t = FLT_MAX;
    best.u = best.v = 0.0f;
    best.leaf = -1;
    best.prim = -1;

    const float3 inv = f3(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z));
    int sp = 1;
    int cur = bvh.n_faces > 1 ? 0 : ~0;   // single-triangle scene: the root is leaf 0
    unsigned nv = 0, nt = 0;
#if TPT_STACK_TOP
    int top = kNoTop;
#define TPT_PUSH(v)                        \
    do {                                   \
        if (top != kNoTop) stack[sp++] = top; \
        top = (v);                         \
    } while (0)
#define TPT_POP(dst)                       \
    do {                                   \
        if (top != kNoTop) {               \
            dst = top;                     \
            top = kNoTop;                  \
        } else {                           \
            dst = stack[--sp];             \
        }                                  \
    } while (0)
#else
#define TPT_PUSH(v) stack[sp++] = (v)
#define TPT_POP(dst) dst = stack[--sp]
#endif

    while (cur != kStackSentinel) {
        while (cur >= 0) {
            const float4* n = bvh.nodes + 4 * (size_t)cur;
            const float4 n0 = __ldg(n), n1 = __ldg(n + 1), n2 = __ldg(n + 2);
            const int2 n3 = __ldg(reinterpret_cast<const int2*>(n + 3));
            if (COUNT) ++nv;
            float loL, loR;
            const bool hL = slab(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, o, inv, best.t, loL);
            const bool hR = slab(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, o, inv, best.t, loR);
            if (hL && hR) {
                const bool rightFirst = loR < loL;
                cur = rightFirst ? n3.y : n3.x;
                TPT_PUSH(rightFirst ? n3.x : n3.y);
            } else if (hL) {
                cur = n3.x;
            } else if (hR) {
                cur = n3.y;
            } else {
                TPT_POP(cur);
            }
        }
        if (cur == kStackSentinel) break;
        if (COUNT) ++nt;
        if (leaf_test(bvh, ~cur, o, d, best) && anyHit) break;
        TPT_POP(cur);
    }
#undef TPT_PUSH
#undef TPT_POP
    if (COUNT) {   // one atomic per converged group of lanes, not per ray
        const unsigned m = __activemask();
        const unsigned sv = __reduce_add_sync(m, nv), st = __reduce_add_sync(m, nt);
        if ((threadIdx.x & 31) == __ffs(m) - 1) {
            atomicAdd(nodeVisits, (unsigned long long)sv);
            atomicAdd(triTests, (unsigned long long)st);
        }
    }
    return best;
}

// ---- traversal as a resumable per-lane state machine (used by the persistent k_wave) -------------------
// cur >= 0: the lane is about to test the two child boxes of internal node `cur`;
// cur <  0 and != kStackSentinel: the lane is about to test triangle ~cur;
// cur == kStackSentinel: traversal finished (best holds the result).
struct TravState {
    float3 o, d, inv;
    HitRec best;
    int cur, sp;
#if TPT_STACK_TOP_SM
    int top;   // newest stack entry, or kNoTop
#endif
};
// stack[0] must hold kStackSentinel (written once by the kernel)
__device__ __forceinline__ void trav_begin(TravState& s, const BvhView& bvh, float3 o, float3 d) {
    s.o = o;
    s.d = d;
    s.inv = f3(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z));
    s.best.t = FLT_MAX;
    s.best.u = s.best.v = 0.0f;
    s.best.leaf = -1;
    s.best.prim = -1;
    s.sp = 1;
#if TPT_STACK_TOP_SM
    s.top = kNoTop;
#endif
    s.cur = bvh.n_faces > 1 ? 0 : ~0;
}
// Traversal stack of the persistent kernels.  A per-thread array in LOCAL memory is interleaved by lane, so a push or pop costs
// one L1 wavefront per DISTINCT stack depth among the lanes of the warp (ncu on k_wave, ball.gltf wave 1: 502 M local
// load/store instructions, 3.5 sectors each — as many data-pipe wavefronts as the node fetches, in a kernel whose limiter is
// exactly that pipe, 84 %).  The first kSmemStack entries therefore live in SHARED memory, entry e of thread t at [e][t]: any
// mix of depths is conflict-free, one wavefront per instruction.  Deeper entries (rare) spill to a local array.  Measured on the
// 10 M-triangle soup (k_trace_rays_persistent): incoherent secondary rays 1 241 -> 1 321 Mrays/s, primary rays unchanged.
#ifndef TPT_SMEM_STACK
#define TPT_SMEM_STACK 16
#endif
constexpr int kSmemStack = TPT_SMEM_STACK;
struct HybridStack {
    int* sh;      // this thread's column of the block's shared array (stride 128 ints); entry 0 holds the sentinel
    int* local;   // entries kSmemStack.. (kStackSize - kSmemStack ints)
    __device__ __forceinline__ void mem_push(int& sp, int v) const {
        if (sp < kSmemStack) sh[sp * 128] = v;
        else local[sp - kSmemStack] = v;
        ++sp;
    }
    __device__ __forceinline__ int mem_pop(int& sp) const {
        --sp;
        return sp < kSmemStack ? sh[sp * 128] : local[sp - kSmemStack];
    }
};
struct LocalStack {
    int* p;
    __device__ __forceinline__ void mem_push(int& sp, int v) const { p[sp++] = v; }
    __device__ __forceinline__ int mem_pop(int& sp) const { return p[--sp]; }
};
template <class Stack>
__device__ __forceinline__ void stack_push(const Stack& st, TravState& s, int v) {
#if TPT_STACK_TOP_SM
    if (s.top != kNoTop) st.mem_push(s.sp, s.top);
    s.top = v;
#else
    st.mem_push(s.sp, v);
#endif
}
template <class Stack>
__device__ __forceinline__ int stack_pop(const Stack& st, TravState& s) {
#if TPT_STACK_TOP_SM
    if (s.top != kNoTop) {
        const int r = s.top;
        s.top = kNoTop;
        return r;
    }
#endif
    return st.mem_pop(s.sp);
}

template <class Stack>
__device__ __forceinline__ void trav_node_step(TravState& s, const BvhView& bvh, const Stack& stack) {
    const float4* n = bvh.nodes + 4 * (size_t)s.cur;
#ifdef TPT_NO_LDG256
    const float4 n0 = __ldg(n), n1 = __ldg(n + 1), n2 = __ldg(n + 2);
    const int2 n3 = __ldg(reinterpret_cast<const int2*>(n + 3));
#else
    const F8 lo8 = ldg256(n), hi8 = ldg256(n + 2);
    const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a;
    const int2 n3 = make_int2(__float_as_int(hi8.b.x), __float_as_int(hi8.b.y));
#endif
    float loL, loR;
    const bool hL = slab(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, s.o, s.inv, s.best.t, loL);
    const bool hR = slab(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, s.o, s.inv, s.best.t, loR);
    if (hL && hR) {
        const bool rightFirst = loR < loL;
        s.cur = rightFirst ? n3.y : n3.x;
        stack_push(stack, s, rightFirst ? n3.x : n3.y);
    } else if (hL) {
        s.cur = n3.x;
    } else if (hR) {
        s.cur = n3.y;
    } else {
        s.cur = stack_pop(stack, s);
    }
}
template <class Stack>
__device__ __forceinline__ void trav_leaf_step(TravState& s, const BvhView& bvh, bool anyHit, const Stack& stack) {
    if (leaf_test(bvh, ~s.cur, s.o, s.d, s.best) && anyHit) s.cur = kStackSentinel;
    else s.cur = stack_pop(stack, s);
}

// ---- wide compressed nodes (BVH4) for trees that do not live in L1 --------------------------------------------------------
// north_star: "compressed BVH nodes ... read with coalesced 128-bit loads".  For incoherent rays on the 10 M-triangle soup the
// BVH2 walk is bound by the L1 data pipe (one wavefront per lane and load instruction, 72.9 node visits per secondary ray) and by
// the chain of dependent node fetches.  A 4-wide node halves both: 64 bytes — two 256-bit loads, like a BVH2 node — hold FOUR child
// boxes quantised to 8 bits per coordinate against the node's own box, so one visit replaces two levels of the binary tree.
//   w0 = (origin.x, origin.y, origin.z, ex | ey << 8 | ez << 16)     box minimum of the node; per-axis scale 2^(e - 127)
//   w1 = child[0..3]            >= 0: internal node (same index as in the binary tree)   < 0: leaf ~child (sorted triangle)
//   w2 = (qlo.x[4], qlo.y[4], qlo.z[4], qhi.x[4])    one byte per child;  w3 = (qhi.y[4], qhi.z[4], -, -)
// A child box decodes as lo = fma(qlo, scale, origin), hi = fma(qhi, scale, origin); the builder (bvh.cu: k_collapse4) rounds qlo
// down and qhi up and CHECKS, with this very fma, that the decoded box contains the exact one, so culling stays conservative
// and (prim, t, u, v) stay bit-identical to the binary walk and to brute force.  An unused slot holds kWideEmpty.  Leaves and their
// exactly-rounded triangle test are unchanged.
constexpr int kWideEmpty = (int)0x80000001;   // neither a node, nor a leaf code (~leaf >= -2^30), nor the stack sentinel
__device__ __forceinline__ float wide_scale(uint32_t e) { return __uint_as_float((e & 255u) << 23); }
// byte c of w as a float.  (Building the float with PRMT + one subtraction instead of shift / mask / I2F was measured SLOWER on the
// 10 M-triangle soup, 1 119 vs 1 274 Mrays/s on the incoherent set: the walk over wide nodes is bound by instruction issue, not by the
// conversion pipe.)
__device__ __forceinline__ float wide_byte(uint32_t w, int c) { return (float)((w >> (8 * c)) & 255u); }

// Ordered traversal pushes at most three children per visit and the tree is at most 95 levels deep (63 key bits + 32 index bits)
constexpr int kWideStackSize = 3 * 96;

template <class Stack>
__device__ __forceinline__ void trav_node_step4(TravState& s, const float4* __restrict__ nodes4, const Stack& stack) {
    const float4* n = nodes4 + 4 * (size_t)s.cur;
    const F8 a = ldg256(n), b = ldg256(n + 2);
    const uint32_t meta = __float_as_uint(a.a.w);
    const float sx = wide_scale(meta), sy = wide_scale(meta >> 8), sz = wide_scale(meta >> 16);
    const uint32_t qlx = __float_as_uint(b.a.x), qly = __float_as_uint(b.a.y), qlz = __float_as_uint(b.a.z);
    const uint32_t qhx = __float_as_uint(b.a.w), qhy = __float_as_uint(b.b.x), qhz = __float_as_uint(b.b.y);
    const int child[4] = {__float_as_int(a.b.x), __float_as_int(a.b.y), __float_as_int(a.b.z), __float_as_int(a.b.w)};
    float t[4];
    int c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float lo;
        const bool hit = slab(fmaf(wide_byte(qlx, k), sx, a.a.x), fmaf(wide_byte(qhx, k), sx, a.a.x), fmaf(wide_byte(qly, k), sy, a.a.y),
                              fmaf(wide_byte(qhy, k), sy, a.a.y), fmaf(wide_byte(qlz, k), sz, a.a.z), fmaf(wide_byte(qhz, k), sz, a.a.z), s.o, s.inv,
                              s.best.t, lo);
        t[k] = (hit && child[k] != kWideEmpty) ? lo : FLT_MAX;
        c[k] = child[k];
    }
    // sort the four (entry distance, child) pairs, nearest first: 5 compare-exchanges; misses (FLT_MAX) sink to the end
#define TPT_CSWAP(i, j)                    \
    if (t[j] < t[i]) {                     \
        const float tt = t[i];             \
        t[i] = t[j];                       \
        t[j] = tt;                         \
        const int cc = c[i];               \
        c[i] = c[j];                       \
        c[j] = cc;                         \
    }
    TPT_CSWAP(0, 1) TPT_CSWAP(2, 3) TPT_CSWAP(0, 2) TPT_CSWAP(1, 3) TPT_CSWAP(1, 2)
#undef TPT_CSWAP
    if (t[0] == FLT_MAX) {
        s.cur = stack_pop(stack, s);
        return;
    }
    if (t[3] != FLT_MAX) stack_push(stack, s, c[3]);   // farthest first: the nearest of the pushed ones pops first
    if (t[2] != FLT_MAX) stack_push(stack, s, c[2]);
    if (t[1] != FLT_MAX) stack_push(stack, s, c[1]);
    s.cur = c[0];
}

// ---- pixel-frustum culling (k_primary) -------------------------------------------------------------------------
// All camera samples of one pixel leave the same origin through the pixel's quadrilateral on the image plane, so
// they lie inside a four-sided pyramid.  If that pyramid misses every leaf box of the tree, none of the pixel's
// samples can hit anything and they need no traversal at all.  Conservative by construction: the footprint is grown
// by 1/64 pixel on each side (rounding of the real ray directions is ~1e-7 of that), a box is rejected only when it
// lies wholly outside one side plane, and a leaf counts as hit as soon as its BOX overlaps the pyramid.
struct PixelFrustum {
    float3 o;
    float3 n[4];   // inward normals of the four side planes (all pass through o)
};
__device__ __forceinline__ bool frustum_box(const PixelFrustum& f, float lox, float hix, float loy, float hiy, float loz, float hiz) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float3 n = f.n[k];
        const float v = n.x * ((n.x >= 0.f ? hix : lox) - f.o.x) + n.y * ((n.y >= 0.f ? hiy : loy) - f.o.y) + n.z * ((n.z >= 0.f ? hiz : loz) - f.o.z);
        if (v < 0.f) return false;   // even the most inward corner of the box is outside this plane
    }
    return true;
}
__device__ __forceinline__ bool frustum_may_hit(const BvhView& bvh, const PixelFrustum& f, int* stack) {
    if (bvh.n_faces <= 1) return true;
    int sp = 1, cur = 0;   // stack[0] holds the sentinel
    while (cur != kStackSentinel) {
        const float4* n = bvh.nodes + 4 * (size_t)cur;
        const float4 n0 = __ldg(n), n1 = __ldg(n + 1), n2 = __ldg(n + 2);
        const int2 n3 = __ldg(reinterpret_cast<const int2*>(n + 3));
        const bool hL = frustum_box(f, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y);
        const bool hR = frustum_box(f, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w);
        if ((hL && n3.x < 0) || (hR && n3.y < 0)) return true;   // a leaf box overlaps the pyramid
        if (hL && hR) {
            stack[sp++] = n3.y;
            cur = n3.x;
        } else if (hL) {
            cur = n3.x;
        } else if (hR) {
            cur = n3.y;
        } else {
            cur = stack[--sp];
        }
    }
    return false;
}

// The same walk, collecting instead of stopping: every leaf whose box overlaps the pyramid is a CANDIDATE, and a
// camera sample of this pixel can only hit candidates.  Testing the candidates directly (leaf_test, the same exactly
// rounded arithmetic and tie-break) gives bit for bit the closest hit a full traversal returns, without walking the
// tree once per sample.  Candidates are kept sorted by a lower bound of the hit distance inside their box (distance
// from the apex to the box, with slack for rounding), so the per-sample loop stops at the first candidate whose bound
// exceeds the best t found.  They live at the TOP of the caller's stack array — leaf k at stack[kStackSize-1-k], its
// bound (float bits) at stack[kStackSize-1-kMaxCandidates-k] — while the walk's own stack grows from the bottom.
// Returns the number of candidates, or -1 when there are more than kMaxCandidates (or the walk would collide with
// them): the caller then traverses per sample.
#ifndef TPT_MAX_CANDIDATES
#define TPT_MAX_CANDIDATES 24
#endif
constexpr int kMaxCandidates = TPT_MAX_CANDIDATES;
__device__ __forceinline__ float box_distance_bound(float3 o, float lox, float hix, float loy, float hiy, float loz, float hiz) {
    const float dx = fmaxf(fmaxf(lox - o.x, o.x - hix), 0.0f), dy = fmaxf(fmaxf(loy - o.y, o.y - hiy), 0.0f),
                dz = fmaxf(fmaxf(loz - o.z, o.z - hiz), 0.0f);
    const float mag = fmaxf(fmaxf(fmaxf(fabsf(lox), fabsf(hix)), fmaxf(fabsf(loy), fabsf(hiy))), fmaxf(fmaxf(fabsf(loz), fabsf(hiz)),
                            fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z))));
    // t is measured in units of |d| <= 1 (the reference's frsqrt never overestimates), so t >= distance; an accepted hit
    // lies within ~1e-6 * mag of its triangle's box: 1e-4 relative + 3e-5 * mag absolute slack
    return fmaxf(sqrtf(dx * dx + dy * dy + dz * dz) * 0.9999f - 3e-5f * mag, 0.0f);
}
__device__ __forceinline__ bool candidate_insert(int* stack, int& nc, int leaf, float bound) {
    if (nc == kMaxCandidates) return false;
    int k = nc++;
    while (k > 0 && __int_as_float(stack[kStackSize - kMaxCandidates - k]) > bound) {   // entry k-1 moves to k
        stack[kStackSize - 1 - k] = stack[kStackSize - k];
        stack[kStackSize - 1 - kMaxCandidates - k] = stack[kStackSize - kMaxCandidates - k];
        --k;
    }
    stack[kStackSize - 1 - k] = leaf;
    stack[kStackSize - 1 - kMaxCandidates - k] = __float_as_int(bound);
    return true;
}
__device__ __forceinline__ int frustum_collect(const BvhView& bvh, const PixelFrustum& f, int* stack, unsigned& nodeVisits) {
    if (bvh.n_faces <= 1) {
        stack[kStackSize - 1] = 0;
        stack[kStackSize - 1 - kMaxCandidates] = 0;   // bound 0.0f
        return 1;
    }
    int sp = 1, cur = 0, nc = 0;   // stack[0] holds the sentinel
    while (cur != kStackSentinel) {
        const float4* n = bvh.nodes + 4 * (size_t)cur;
        const float4 n0 = __ldg(n), n1 = __ldg(n + 1), n2 = __ldg(n + 2);
        const int2 n3 = __ldg(reinterpret_cast<const int2*>(n + 3));
        ++nodeVisits;
        bool hL = frustum_box(f, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y);
        bool hR = frustum_box(f, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w);
        if (hL && n3.x < 0) {
            if (!candidate_insert(stack, nc, ~n3.x, box_distance_bound(f.o, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y))) return -1;
            hL = false;
        }
        if (hR && n3.y < 0) {
            if (!candidate_insert(stack, nc, ~n3.y, box_distance_bound(f.o, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w))) return -1;
            hR = false;
        }
        if (hL && hR) {
            if (sp >= kStackSize - 2 * kMaxCandidates - 1) return -1;
            stack[sp++] = n3.y;
            cur = n3.x;
        } else if (hL) {
            cur = n3.x;
        } else if (hR) {
            cur = n3.y;
        } else {
            cur = stack[--sp];
        }
    }
    return nc;
}

}   // namespace tpt
