// bvh.cu — parallel LBVH build for sm_100a.
//
// Replaces the `transform` kernel (reference src/path_tracer.cu:227-263) and BVH::construct
// (src/bvh.cu:128-331).  Pipeline, all on the context stream with no host synchronisation:
//   1. k_transform      per vertex: object->world position and (frsqrt-normalised) normal, and the
//                       scene AABB by warp shuffle + one ordered-uint atomicMin/Max per warp.
//   2. k_morton         per face: AABB centre -> 63-bit Morton key normalised to the SCENE bounds
//                       (the reference uses a fixed 2^-12 grid over [-256,256) and saturates outside it).
//   3. radix sort       (key64, face32) pairs — the reference sorts whole 36-byte nodes.
//   4. k_hierarchy      Karras 2012 ranges/splits with (key, index) tie-break, so any number of equal
//                       keys gives a valid tree (the reference breaks on runs of >= 3).
//   5. k_leaves_refit   per sorted leaf: write the 48-B triangle and normal records, then walk to the
//                       root with one atomic arrival counter per node; the second arrival owns the
//                       node.  Grid-wide, replaces the reference's single 1024-thread block
//                       (computeBBox, src/bvh.cu:220-302).
#include <algorithm>
#include <atomic>

#include "scene.cuh"
#include "sort.cuh"
#include "traverse.cuh"

namespace tpt {

__global__ void k_init_bounds(uint32_t* bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 8) bounds[threadIdx.x] = 0u;   // [3..5] max, [6] refit frontier count, [7] spare
}

// Built once per scene: which object each vertex belongs to, and the material of each face
// (replaces findIdxOfTrans, src/path_tracer.cu:227-237, and mtlLinearSearch, :125-135, both O(#objects)
// scans per use in the reference).
__global__ void k_face_tables(const uint32_t* __restrict__ indices, const tpt_mtl_interval* __restrict__ objects, int nObjects,
                              uint32_t nFaces, uint32_t* __restrict__ vertObj, uint32_t* __restrict__ faceMtl) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nFaces) return;
    // last object whose begin <= f (objects are in ascending begin order)
    int lo = 0, hi = nObjects - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((uint32_t)objects[mid].begin <= f) lo = mid;
        else hi = mid - 1;
    }
    faceMtl[f] = (uint32_t)objects[lo].mtl_idx;
    vertObj[indices[3 * f + 0]] = (uint32_t)lo;
    vertObj[indices[3 * f + 1]] = (uint32_t)lo;
    vertObj[indices[3 * f + 2]] = (uint32_t)lo;
}

// NORMALS = false (scenes whose normal records are made on first render, tpt_scene_ensure_normals): positions and bounds only —
// a build that is only traced never reads a normal; k_transform_normals catches up when a render needs them.
template <bool NORMALS>
__global__ void __launch_bounds__(256) k_transform(const float* __restrict__ objVerts, const float* __restrict__ objNorms,
                                                   const uint32_t* __restrict__ vertObj, const float* __restrict__ vertTrans,
                                                   const float* __restrict__ normalTrans, uint32_t nVerts,
                                                   float* __restrict__ worldVerts, float* __restrict__ worldNorms,
                                                   uint32_t* __restrict__ bounds) {
    float3 mn = f3(FLT_MAX, FLT_MAX, FLT_MAX), mx = f3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nVerts; v += gridDim.x * blockDim.x) {
        const uint32_t o = vertObj[v];
        if (o == 0xffffffffu) continue;   // not referenced by any face: the reference never writes it either
        const float3 p = f3(objVerts[3 * v], objVerts[3 * v + 1], objVerts[3 * v + 2]);
        const float3 wp = xmat4_point(vertTrans + 16 * (size_t)o, p);
        worldVerts[3 * v] = wp.x; worldVerts[3 * v + 1] = wp.y; worldVerts[3 * v + 2] = wp.z;
        if (NORMALS) {
            const float3 n = f3(objNorms[3 * v], objNorms[3 * v + 1], objNorms[3 * v + 2]);
            const float3 wn = xnormalize(xmat4_dir(normalTrans + 16 * (size_t)o, n));
            worldNorms[3 * v] = wn.x; worldNorms[3 * v + 1] = wn.y; worldNorms[3 * v + 2] = wn.z;
        }
        mn = fmin3(mn, wp);
        mx = fmax3(mx, wp);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        mn.x = fminf(mn.x, __shfl_xor_sync(0xffffffffu, mn.x, off));
        mn.y = fminf(mn.y, __shfl_xor_sync(0xffffffffu, mn.y, off));
        mn.z = fminf(mn.z, __shfl_xor_sync(0xffffffffu, mn.z, off));
        mx.x = fmaxf(mx.x, __shfl_xor_sync(0xffffffffu, mx.x, off));
        mx.y = fmaxf(mx.y, __shfl_xor_sync(0xffffffffu, mx.y, off));
        mx.z = fmaxf(mx.z, __shfl_xor_sync(0xffffffffu, mx.z, off));
    }
    if ((threadIdx.x & 31) == 0 && mn.x <= mx.x) {
        atomicMin(bounds + 0, f2ord(mn.x)); atomicMin(bounds + 1, f2ord(mn.y)); atomicMin(bounds + 2, f2ord(mn.z));
        atomicMax(bounds + 3, f2ord(mx.x)); atomicMax(bounds + 4, f2ord(mx.y)); atomicMax(bounds + 5, f2ord(mx.z));
    }
}

__global__ void __launch_bounds__(256) k_transform_normals(const float* __restrict__ objNorms, const uint32_t* __restrict__ vertObj,
                                                           const float* __restrict__ normalTrans, uint32_t nVerts, float* __restrict__ worldNorms) {
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nVerts; v += gridDim.x * blockDim.x) {
        const uint32_t o = vertObj[v];
        if (o == 0xffffffffu) continue;
        const float3 n = f3(objNorms[3 * v], objNorms[3 * v + 1], objNorms[3 * v + 2]);
        const float3 wn = xnormalize(xmat4_dir(normalTrans + 16 * (size_t)o, n));   // the same arithmetic as k_transform<true>
        worldNorms[3 * v] = wn.x; worldNorms[3 * v + 1] = wn.y; worldNorms[3 * v + 2] = wn.z;
    }
}

// Common-prefix length of the augmented keys (key, index): equal keys fall back to the index bits.
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a != b) return __clzll((long long)(a ^ b));
    return 64 + __clz(i ^ j);
}

__global__ void __launch_bounds__(256) k_hierarchy(const unsigned long long* __restrict__ keys, int nFaces, float4* __restrict__ nodes,
                                                   int* __restrict__ parentInt, int* __restrict__ parentLeaf, int* __restrict__ rangeOther) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nFaces - 1) return;
    const int d = delta(keys, nFaces, i, i + 1) > delta(keys, nFaces, i, i - 1) ? 1 : -1;
    const int dMin = delta(keys, nFaces, i, i - d);
    int lmax = 2;
    while (delta(keys, nFaces, i, i + lmax * d) > dMin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t > 0; t >>= 1)
        if (delta(keys, nFaces, i, i + (l + t) * d) > dMin) l += t;
    const int j = i + l * d;
    rangeOther[i] = j;   // node i covers the sorted leaves [min(i,j), max(i,j)]
    const int dNode = delta(keys, nFaces, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, nFaces, i, i + (s + t) * d) > dNode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? ~gamma : gamma;
    const int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    if (left < 0) parentLeaf[~left] = i; else parentInt[left] = i;
    if (right < 0) parentLeaf[~right] = i; else parentInt[right] = i;
    if (i == 0) parentInt[0] = -1;
    reinterpret_cast<int4*>(nodes)[4 * (size_t)i + 3] = make_int4(left, right, 0, 0);
}

__device__ __forceinline__ void store_child_box(float4* node, bool isLeft, float3 mn, float3 mx) {
    // n0/n1 hold x and y of the left/right child, n2 holds z of both (see scene.cuh).
    float* n = reinterpret_cast<float*>(node);
    float4* xy = node + (isLeft ? 0 : 1);
    __stcg(xy, make_float4(mn.x, mx.x, mn.y, mx.y));
    __stcg(reinterpret_cast<float2*>(n + 8 + (isLeft ? 0 : 2)), make_float2(mn.z, mx.z));
}

// Refit, two kernels.  A bottom-up walk with one global atomic + two fences per level keeps a 256-thread block
// resident for as long as its longest chain, and at 10 M triangles that residency — not bandwidth — set the
// build time.  So:
//   k_leaves_refit  (one thread per sorted leaf) writes the triangle/normal records and walks up while the
//                   node's whole leaf range lies inside this block's 256 leaves: sibling boxes are exchanged
//                   through shared memory (shared atomics, block-scope fences).  At the first node that spans
//                   blocks it falls back to the global arrival counter; the second arrival does not continue
//                   but appends the (now complete) node to a frontier list.
//   k_refit_upper   one thread per frontier node finishes the walk to the root with the global protocol; there
//                   are only ~F/100 of them, all resident at once.
constexpr int kRefitBlock = 256;

__global__ void __launch_bounds__(kRefitBlock) k_leaves_refit(const uint32_t* __restrict__ sortedFaces, const uint32_t* __restrict__ indices,
                                                      const float* __restrict__ worldVerts, const float* __restrict__ worldNorms,
                                                      const uint32_t* __restrict__ faceMtl, int nFaces, const int* __restrict__ parentInt,
                                                      const int* __restrict__ parentLeaf, const int* __restrict__ rangeOther,
                                                      uint32_t* __restrict__ flags, float4* __restrict__ nodes, float4* __restrict__ triGeom,
                                                      float4* __restrict__ triNorm, int* __restrict__ frontier, uint32_t* __restrict__ frontierCount) {
    __shared__ unsigned sFlag[kRefitBlock];
    __shared__ float sBox[kRefitBlock][2][6];
    sFlag[threadIdx.x] = 0u;
    __syncthreads();
    const int b0 = blockIdx.x * kRefitBlock;
    const int leaf = b0 + threadIdx.x;
    if (leaf >= nFaces) return;
    const uint32_t f = sortedFaces[leaf];
    const uint32_t i0 = indices[3 * f], i1 = indices[3 * f + 1], i2 = indices[3 * f + 2];
    const float3 a = f3(worldVerts[3 * i0], worldVerts[3 * i0 + 1], worldVerts[3 * i0 + 2]);
    const float3 b = f3(worldVerts[3 * i1], worldVerts[3 * i1 + 1], worldVerts[3 * i1 + 2]);
    const float3 c = f3(worldVerts[3 * i2], worldVerts[3 * i2 + 1], worldVerts[3 * i2 + 2]);
    const float3 e1 = xsub3(b, a), e2 = xsub3(c, a);   // exactly the edges Möller–Trumbore forms per test in the reference
    __stcs(triGeom + 3 * (size_t)leaf + 0, make_float4(a.x, a.y, a.z, e1.x));
    __stcs(triGeom + 3 * (size_t)leaf + 1, make_float4(e1.y, e1.z, e2.x, e2.y));
    __stcs(triGeom + 3 * (size_t)leaf + 2, make_float4(e2.z, __int_as_float((int)f), __int_as_float((int)faceMtl[f]), 0.f));
    if (triNorm != nullptr) {   // small scenes: with the leaf; large ones: k_leaf_normals, when a render first needs them
        const float3 n0 = f3(worldNorms[3 * i0], worldNorms[3 * i0 + 1], worldNorms[3 * i0 + 2]);
        const float3 n1 = f3(worldNorms[3 * i1], worldNorms[3 * i1 + 1], worldNorms[3 * i1 + 2]);
        const float3 n2 = f3(worldNorms[3 * i2], worldNorms[3 * i2 + 1], worldNorms[3 * i2 + 2]);
        __stcs(triNorm + 3 * (size_t)leaf + 0, make_float4(n0.x, n0.y, n0.z, n1.x));
        __stcs(triNorm + 3 * (size_t)leaf + 1, make_float4(n1.y, n1.z, n2.x, n2.y));
        __stcs(triNorm + 3 * (size_t)leaf + 2, make_float4(n2.z, 0.f, 0.f, 0.f));
    }
    if (nFaces == 1) return;

    float3 mn = fmin3(fmin3(a, b), c), mx = fmax3(fmax3(a, b), c);
    int child = ~leaf;
    int cur = parentLeaf[leaf];
    while (cur >= 0) {
        float4* node = nodes + 4 * (size_t)cur;
        const int2 ch = *reinterpret_cast<const int2*>(node + 3);
        const bool isLeft = (ch.x == child);
        store_child_box(node, isLeft, mn, mx);
        const int other = rangeOther[cur];
        const int lo = min(cur, other), hi = max(cur, other);
        if (lo >= b0 && hi < b0 + kRefitBlock) {
            // both children arrive from this block: exchange through shared memory
            const int k = cur - b0, side = isLeft ? 0 : 1;
            float* mine = sBox[k][side];
            mine[0] = mn.x; mine[1] = mn.y; mine[2] = mn.z; mine[3] = mx.x; mine[4] = mx.y; mine[5] = mx.z;
            __threadfence_block();
            if (atomicAdd(&sFlag[k], 1u) == 0u) return;   // first arrival: the sibling will finish this node
            __threadfence_block();
            const volatile float* sib = sBox[k][side ^ 1];
            mn = fmin3(mn, f3(sib[0], sib[1], sib[2]));
            mx = fmax3(mx, f3(sib[3], sib[4], sib[5]));
            child = cur;
            cur = parentInt[cur];
        } else {
            __threadfence();
            if (atomicAdd(flags + cur, 1u) == 1u) frontier[atomicAdd(frontierCount, 1u)] = cur;   // complete: k_refit_upper goes on
            return;
        }
    }
}

// ---- hierarchy emitted BY the bottom-up walk (full builds) ---------------------------------------------------------------------
// k_hierarchy finds every node's range and split with binary searches over the keys — as many steps as the longest range in the
// warp, 14 of 32 lanes busy, 0.29 ms at 10 M triangles — and only then can the refit walk start.  The walk can find the tree
// itself (Apetrei 2014): a finished node [l, r] joins the neighbour it shares the longer key prefix with — to the right if
// delta(r, r+1) > delta(l-1, l) (it is the LEFT child of the node split between r and r+1), else to the left — and meets its
// sibling at that split.  Two key reads per level instead of a search, no separate pass.  The numbering stays Karras': a left
// child is node `r` (its range's right end), a right child node `l`, the root node 0 — so the nodes, parents and range ends
// written here are THE arrays k_hierarchy + k_leaves_refit + k_refit_upper produce, bit for bit (tests/test_gpu_parity.py
// compares them with the reference's tree; TPT_BUILD_KARRAS=1 runs the old kernels).  What changes is where siblings meet: a node's
// own index is only known once BOTH children are there (its direction depends on its whole range), so the first arrival cannot
// write into the node; it leaves (box, id, far end of its range) in a slot of the SPLIT position instead, and the second arrival
// writes the finished 64-byte node in one piece.
//   k_leaves_build  one thread per sorted leaf: triangle record, then up while the parent's whole range lies inside the block's
//                   256 leaves (slots and keys in shared memory).  Whether it does is one more prefix comparison: the parent of
//                   a left child ends before the block does iff delta(split+1, first key after the block) < delta(split, split+1).
//                   At the first parent that spans blocks: global slot + arrival counter; the second arrival queues the split.
//   k_build_upper   one thread per queued split finishes the walk to the root with the global protocol (as k_refit_upper).
struct MergeHalf {   // 32 bytes: what a child leaves at its parent's split
    float mnx, mny, mnz, mxx, mxy, mxz;
    int id;    // ~leaf, or the child's node index
    int end;   // the far end of the child's range: its first leaf for a left child, its last for a right child
};
__device__ __forceinline__ void merge_store_global(float4* slot, const MergeHalf& h) {
    __stcg(slot, make_float4(h.mnx, h.mny, h.mnz, h.mxx));
    __stcg(slot + 1, make_float4(h.mxy, h.mxz, __int_as_float(h.id), __int_as_float(h.end)));
}
__device__ __forceinline__ MergeHalf merge_load_global(const float4* slot) {
    const float4 a = __ldcg(slot), b = __ldcg(slot + 1);
    return MergeHalf{a.x, a.y, a.z, a.w, b.x, b.y, __float_as_int(b.z), __float_as_int(b.w)};
}
// the finished node `k` with children L (range [first, split]) and R ([split+1, last]): record, parents, range end
__device__ __forceinline__ void write_node(float4* __restrict__ nodes, int* __restrict__ parentInt, int* __restrict__ parentLeaf, int* __restrict__ rangeOther,
                                           int k, const MergeHalf& L, const MergeHalf& R) {
    float4* node = nodes + 4 * (size_t)k;
    __stcg(node + 0, make_float4(L.mnx, L.mxx, L.mny, L.mxy));
    __stcg(node + 1, make_float4(R.mnx, R.mxx, R.mny, R.mxy));
    __stcg(node + 2, make_float4(L.mnz, L.mxz, R.mnz, R.mxz));
    __stcg(reinterpret_cast<int4*>(node + 3), make_int4(L.id, R.id, 0, 0));
    if (L.id < 0) parentLeaf[~L.id] = k; else parentInt[L.id] = k;
    if (R.id < 0) parentLeaf[~R.id] = k; else parentInt[R.id] = k;
    rangeOther[k] = (k == R.end) ? L.end : R.end;   // node k is one end of its range [L.end, R.end]; the root (k = 0 = L.end) gets R.end
}
__device__ __forceinline__ MergeHalf merge_union(const MergeHalf& L, const MergeHalf& R) {
    return MergeHalf{fminf(L.mnx, R.mnx), fminf(L.mny, R.mny), fminf(L.mnz, R.mnz), fmaxf(L.mxx, R.mxx), fmaxf(L.mxy, R.mxy), fmaxf(L.mxz, R.mxz), 0, 0};
}

// The walk inside a block needs delta() only between NEIGHBOURS and between a position and the block's edges, and for sorted
// (key, index) pairs delta(x, y) is the minimum of the neighbour deltas in between.  So the block computes adj[j] = delta(j, j+1)
// for j = b0-1 .. bEnd once (one coalesced key read per thread) and its running minima from either edge; a step of the walk is
// then three shared-memory integer loads instead of six 64-bit key comparisons.
constexpr int kDeltaInf = 1 << 20;
__global__ void __launch_bounds__(kRefitBlock) k_leaves_build(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ sortedFaces,
                                                      const uint32_t* __restrict__ indices, const float* __restrict__ worldVerts,
                                                      const float* __restrict__ worldNorms, const uint32_t* __restrict__ faceMtl, int nFaces,
                                                      int* __restrict__ parentInt, int* __restrict__ parentLeaf, int* __restrict__ rangeOther,
                                                      uint32_t* __restrict__ flags, float4* __restrict__ slots, float4* __restrict__ nodes,
                                                      float4* __restrict__ triGeom, float4* __restrict__ triNorm, int* __restrict__ frontier,
                                                      uint32_t* __restrict__ frontierCount) {
    __shared__ unsigned sFlag[kRefitBlock];
    __shared__ float4 sHalf[kRefitBlock][2][2];   // [split - b0][side]: a MergeHalf as two 128-bit words
    __shared__ int sAdj[kRefitBlock + 1];    // sAdj[i] = delta(w0 + i, w0 + i + 1), w0 = b0 - 1: the pairs (b0-1, b0) ... (bEnd, bEnd+1); -1 beyond the array
    __shared__ int sPre[kRefitBlock + 1];    // sPre[i + 1] = min(sAdj[0 .. i]) = delta(b0 - 1, w0 + i + 1); sPre[0] = inf (empty)
    __shared__ int sSuf[kRefitBlock + 2];    // sSuf[i] = min(sAdj[i .. 256]) = delta(w0 + i, bEnd + 1), i >= 1; sSuf[257] = inf (empty)
    __shared__ int sWarp[2][kRefitBlock / 32];
    const int b0 = blockIdx.x * kRefitBlock, w0 = b0 - 1;   // the block's leaves are b0 .. bEnd = b0 + 255
    const int t = threadIdx.x, lane = t & 31, wrp = t >> 5;
    sFlag[t] = 0u;
    {
        const int a0 = delta(keys, nFaces, min(max(w0 + t, 0), nFaces - 1), (w0 + t < 0 || w0 + t >= nFaces) ? -1 : w0 + t + 1);         // pair (w0+t, w0+t+1)
        const int a1 = delta(keys, nFaces, min(w0 + t + 1, nFaces - 1), (w0 + t + 1 >= nFaces) ? -1 : w0 + t + 2);                       // pair (w0+t+1, w0+t+2)
        sAdj[t] = a0;
        if (t == kRefitBlock - 1) sAdj[kRefitBlock] = a1;
        // running minima: prefix over sAdj[0..255] (thread t owns element t), suffix over sAdj[1..256] (thread t owns element t + 1 = a1)
        int pre = a0, suf = a1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int x = __shfl_up_sync(0xffffffffu, pre, o), y = __shfl_down_sync(0xffffffffu, suf, o);
            if (lane >= o) pre = min(pre, x);
            if (lane + o < 32) suf = min(suf, y);
        }
        if (lane == 31) sWarp[0][wrp] = pre;
        if (lane == 0) sWarp[1][wrp] = suf;
        __syncthreads();
        for (int w = 0; w < wrp; ++w) pre = min(pre, sWarp[0][w]);
        for (int w = wrp + 1; w < kRefitBlock / 32; ++w) suf = min(suf, sWarp[1][w]);
        sPre[t + 1] = pre;
        sSuf[t + 1] = suf;
        if (t == 0) {
            sPre[0] = kDeltaInf;
            sSuf[kRefitBlock + 1] = kDeltaInf;
            sSuf[0] = kDeltaInf;   // never read
        }
    }
    __syncthreads();
    const int leaf = b0 + t;
    if (leaf >= nFaces) return;
    const uint32_t f = sortedFaces[leaf];
    const uint32_t i0 = indices[3 * f], i1 = indices[3 * f + 1], i2 = indices[3 * f + 2];
    const float3 a = f3(worldVerts[3 * i0], worldVerts[3 * i0 + 1], worldVerts[3 * i0 + 2]);
    const float3 b = f3(worldVerts[3 * i1], worldVerts[3 * i1 + 1], worldVerts[3 * i1 + 2]);
    const float3 c = f3(worldVerts[3 * i2], worldVerts[3 * i2 + 1], worldVerts[3 * i2 + 2]);
    const float3 e1 = xsub3(b, a), e2 = xsub3(c, a);   // exactly the edges Möller–Trumbore forms per test in the reference
    __stcs(triGeom + 3 * (size_t)leaf + 0, make_float4(a.x, a.y, a.z, e1.x));
    __stcs(triGeom + 3 * (size_t)leaf + 1, make_float4(e1.y, e1.z, e2.x, e2.y));
    __stcs(triGeom + 3 * (size_t)leaf + 2, make_float4(e2.z, __int_as_float((int)f), __int_as_float((int)faceMtl[f]), 0.f));
    if (triNorm != nullptr) {   // small scenes: with the leaf; large ones: k_leaf_normals, when a render first needs them
        const float3 n0 = f3(worldNorms[3 * i0], worldNorms[3 * i0 + 1], worldNorms[3 * i0 + 2]);
        const float3 n1 = f3(worldNorms[3 * i1], worldNorms[3 * i1 + 1], worldNorms[3 * i1 + 2]);
        const float3 n2 = f3(worldNorms[3 * i2], worldNorms[3 * i2 + 1], worldNorms[3 * i2 + 2]);
        __stcs(triNorm + 3 * (size_t)leaf + 0, make_float4(n0.x, n0.y, n0.z, n1.x));
        __stcs(triNorm + 3 * (size_t)leaf + 1, make_float4(n1.y, n1.z, n2.x, n2.y));
        __stcs(triNorm + 3 * (size_t)leaf + 2, make_float4(n2.z, 0.f, 0.f, 0.f));
    }
    if (nFaces == 1) return;

    float3 mn = fmin3(fmin3(a, b), c), mx = fmax3(fmax3(a, b), c);
    int id = ~leaf, l = leaf, r = leaf;
    // which way the current node goes: the LEFT child of the node split at r (delta(r, r+1) is the longer prefix), or the RIGHT
    // child of the node split at l - 1.  adj(j) = sAdj[j - w0]; the ends of the array have adj = -1, so they decide themselves.
    bool goRight = sAdj[r - w0] > sAdj[l - 1 - w0];
    while (true) {
        const int split = goRight ? r : l - 1, side = goRight ? 0 : 1;
        const int end = goRight ? l : r;   // the far end of this node's range
        const int q = sAdj[split - w0];
        // does the parent's whole range lie in this block?  left child: iff delta(split + 1, bEnd + 1) < q; right child: iff
        // delta(b0 - 1, split) < q (empty ranges — the sibling starts outside the block — are inf: no)
        const bool inBlock = goRight ? sSuf[split + 1 - w0] < q : sPre[split - w0] < q;
        const float4 h0 = make_float4(mn.x, mn.y, mn.z, mx.x), h1 = make_float4(mx.y, mx.z, __int_as_float(id), __int_as_float(end));
        if (!inBlock) {
            float4* slot = slots + 4 * (size_t)split + 2 * side;
            __stcg(slot, h0);
            __stcg(slot + 1, h1);
            __threadfence();
            if (atomicAdd(flags + split, 1u) == 1u) frontier[atomicAdd(frontierCount, 1u)] = split;   // complete: k_build_upper goes on
            return;
        }
        const int ks = split - b0;
        sHalf[ks][side][0] = h0;
        sHalf[ks][side][1] = h1;
        __threadfence_block();
        if (atomicAdd(&sFlag[ks], 1u) == 0u) return;   // first arrival: the sibling will finish this node
        __threadfence_block();
        float4 s0, s1;
        {   // the sibling's half, written by another thread of the block: volatile 128-bit shared loads
            const unsigned addr = (unsigned)__cvta_generic_to_shared(&sHalf[ks][side ^ 1][0]);
            asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(s0.x), "=f"(s0.y), "=f"(s0.z), "=f"(s0.w) : "r"(addr));
            asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(s1.x), "=f"(s1.y), "=f"(s1.z), "=f"(s1.w) : "r"(addr + 16u));
        }
        const float3 smn = f3(s0.x, s0.y, s0.z), smx = f3(s0.w, s1.x, s1.y);
        const int sid = __float_as_int(s1.z), send = __float_as_int(s1.w);
        l = goRight ? end : send;   // parent's range: the left child's first leaf ... the right child's last
        r = goRight ? send : end;
        int k = 0;
        const bool root = l == 0 && r == nFaces - 1;
        bool up = false;
        if (!root) {
            up = sAdj[r - w0] > sAdj[l - 1 - w0];
            k = up ? r : l;
        }
        // the finished node k: each child's box and id go to its side of the record (layout: scene.cuh), no data is selected
        float4* node = nodes + 4 * (size_t)k;
        store_child_box(node, side == 0, mn, mx);
        store_child_box(node, side != 0, smn, smx);
        int* ids = reinterpret_cast<int*>(node + 3);
        __stcg(ids + side, id);
        __stcg(ids + (side ^ 1), sid);
        __stcg(reinterpret_cast<int2*>(ids + 2), make_int2(0, 0));
        if (id < 0) parentLeaf[~id] = k; else parentInt[id] = k;
        if (sid < 0) parentLeaf[~sid] = k; else parentInt[sid] = k;
        rangeOther[k] = (k == r) ? l : r;   // node k is one end of its range; the root (k = 0 = l) gets r
        if (root) {
            parentInt[0] = -1;
            return;
        }
        mn = fmin3(mn, smn);
        mx = fmax3(mx, smx);
        id = k;
        goRight = up;
    }
}

__global__ void __launch_bounds__(256) k_build_upper(const unsigned long long* __restrict__ keys, int nFaces, const int* __restrict__ frontier,
                                                     const uint32_t* __restrict__ frontierCount, int* __restrict__ parentInt,
                                                     int* __restrict__ parentLeaf, int* __restrict__ rangeOther, uint32_t* __restrict__ flags,
                                                     float4* __restrict__ slots, float4* __restrict__ nodes) {
    // grid-stride over the queued splits: there are ~F/100 of them (F/2 only for a degenerate chain of a tree), so the launch is
    // sized for the usual case and loops in the other
    const uint32_t count = *frontierCount;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < count; t += gridDim.x * blockDim.x) {
    int split = frontier[t];
    MergeHalf L = merge_load_global(slots + 4 * (size_t)split), R = merge_load_global(slots + 4 * (size_t)split + 2);
    while (true) {
        const int l = L.end, r = R.end;
        const bool root = l == 0 && r == nFaces - 1;
        bool goRight = false;
        int k = 0;
        if (!root) {
            goRight = l == 0 ? true : r == nFaces - 1 ? false : delta(keys, nFaces, r, r + 1) > delta(keys, nFaces, l - 1, l);
            k = goRight ? r : l;
        }
        write_node(nodes, parentInt, parentLeaf, rangeOther, k, L, R);
        if (root) {
            parentInt[0] = -1;
            break;
        }
        MergeHalf me = merge_union(L, R);
        me.id = k;
        me.end = goRight ? l : r;
        split = goRight ? r : l - 1;
        const int side = goRight ? 0 : 1;
        float4* slot = slots + 4 * (size_t)split;
        merge_store_global(slot + 2 * side, me);
        __threadfence();
        if (atomicAdd(flags + split, 1u) == 0u) break;   // first arrival: the sibling will finish this node
        __threadfence();
        const MergeHalf sib = merge_load_global(slot + 2 * (side ^ 1));   // written by another SM: L2 loads
        if (side == 0) {
            L = me;
            R = sib;
        } else {
            L = sib;
            R = me;
        }
    }
    }
}

// The per-leaf shading-normal records (read only by the renderer's shade step, never by traversal), for scenes whose build
// left them out: a 10 M-triangle build that is only traced (tpt_trace_rays) does not pay 36 B of gathers + 48 B of writes per leaf.
__global__ void __launch_bounds__(256) k_leaf_normals(const uint32_t* __restrict__ sortedFaces, const uint32_t* __restrict__ indices,
                                                      const float* __restrict__ worldNorms, int nFaces, float4* __restrict__ triNorm) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= nFaces) return;
    const uint32_t f = sortedFaces[leaf];
    const uint32_t i0 = indices[3 * f], i1 = indices[3 * f + 1], i2 = indices[3 * f + 2];
    const float3 n0 = f3(worldNorms[3 * i0], worldNorms[3 * i0 + 1], worldNorms[3 * i0 + 2]);
    const float3 n1 = f3(worldNorms[3 * i1], worldNorms[3 * i1 + 1], worldNorms[3 * i1 + 2]);
    const float3 n2 = f3(worldNorms[3 * i2], worldNorms[3 * i2 + 1], worldNorms[3 * i2 + 2]);
    __stcs(triNorm + 3 * (size_t)leaf + 0, make_float4(n0.x, n0.y, n0.z, n1.x));
    __stcs(triNorm + 3 * (size_t)leaf + 1, make_float4(n1.y, n1.z, n2.x, n2.y));
    __stcs(triNorm + 3 * (size_t)leaf + 2, make_float4(n2.z, 0.f, 0.f, 0.f));
}

__global__ void __launch_bounds__(256) k_refit_upper(const int* __restrict__ frontier, const uint32_t* __restrict__ frontierCount,
                                                     const int* __restrict__ parentInt, uint32_t* __restrict__ flags,
                                                     float4* __restrict__ nodes) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *frontierCount) return;
    int child = frontier[t];
    float3 mn, mx;
    {
        const float4* node = nodes + 4 * (size_t)child;
        const float4 n0 = __ldcg(node), n1 = __ldcg(node + 1), n2 = __ldcg(node + 2);
        mn = f3(fminf(n0.x, n1.x), fminf(n0.z, n1.z), fminf(n2.x, n2.z));
        mx = f3(fmaxf(n0.y, n1.y), fmaxf(n0.w, n1.w), fmaxf(n2.y, n2.w));
    }
    int cur = parentInt[child];
    while (cur >= 0) {
        float4* node = nodes + 4 * (size_t)cur;
        const int2 ch = *reinterpret_cast<const int2*>(node + 3);
        const bool isLeft = (ch.x == child);
        store_child_box(node, isLeft, mn, mx);
        __threadfence();
        if (atomicAdd(flags + cur, 1u) == 0u) return;   // first arrival: the sibling will finish this node
        __threadfence();
        // second arrival: read the sibling's box (written by another SM: bypass L1) and go up
        const float4 sxy = __ldcg(node + (isLeft ? 1 : 0));
        const float2 sz = __ldcg(reinterpret_cast<const float2*>(reinterpret_cast<const float*>(node) + 8 + (isLeft ? 2 : 0)));
        mn = fmin3(mn, f3(sxy.x, sxy.z, sz.x));
        mx = fmax3(mx, f3(sxy.y, sxy.w, sz.y));
        child = cur;
        cur = parentInt[cur];
    }
}

// ---- 4-wide compressed nodes (layout and decode: traverse.cuh "wide compressed nodes") ------------------------------------------
// One thread per binary node i writes wide node i: it starts from i's two children and twice replaces the internal child with
// the largest surface area by that child's own two children (greedy surface-area collapse: a node absorbs up to three binary
// nodes and has 2..4 children).  EVERY binary node gets a wide node at its own index, reachable or not (about half are): no
// top-down pass, no allocation, no queue — the kernel reads at most three 64-B nodes and writes one.  Child boxes are quantised to
// 8 bits per coordinate against the node's box (= the union of its children, exactly), lower bounds rounded down and upper
// bounds up, and each is re-decoded with the traversal's own fma and corrected until the decoded box contains the exact one.
struct WChild {
    int id;
    float3 lo, hi;
};
__device__ __forceinline__ void wide_children(const float4* __restrict__ nodes, int i, WChild& l, WChild& r) {
    const float4* n = nodes + 4 * (size_t)i;
    const float4 n0 = n[0], n1 = n[1], n2 = n[2];
    const int2 ch = *reinterpret_cast<const int2*>(n + 3);
    l.id = ch.x; l.lo = f3(n0.x, n0.z, n2.x); l.hi = f3(n0.y, n0.w, n2.y);
    r.id = ch.y; r.lo = f3(n1.x, n1.z, n2.z); r.hi = f3(n1.y, n1.w, n2.w);
}
__device__ __forceinline__ float wide_area(const WChild& c) {
    const float3 e = c.hi - c.lo;
    return e.x * e.y + e.y * e.z + e.z * e.x;
}
// smallest biased exponent eb (scale 2^(eb-127)) whose 255 steps from lo reach hi under the decode arithmetic
__device__ __forceinline__ uint32_t wide_exponent(float lo, float hi) {
    int e = -126;
    const float ext = hi - lo;
    if (ext > 0.0f) {
        (void)frexpf(ext * (1.0f / 255.0f), &e);   // ext / 255 = m * 2^e, m in [0.5, 1)
        e = max(-126, min(e, 127));
    }
    while (e < 127 && fmaf(255.0f, __uint_as_float((uint32_t)(e + 127) << 23), lo) < hi) ++e;
    return (uint32_t)(e + 127);
}
__device__ __forceinline__ uint32_t wide_qlo(float v, float origin, float scale, float invScale) {
    int q = (int)fminf(fmaxf(floorf((v - origin) * invScale), 0.0f), 255.0f);
    while (q > 0 && fmaf((float)q, scale, origin) > v) --q;
    return (uint32_t)q;
}
__device__ __forceinline__ uint32_t wide_qhi(float v, float origin, float scale, float invScale) {
    int q = (int)fminf(fmaxf(ceilf((v - origin) * invScale), 0.0f), 255.0f);
    while (q < 255 && fmaf((float)q, scale, origin) < v) ++q;
    return (uint32_t)q;
}

__global__ void __launch_bounds__(256) k_collapse4(const float4* __restrict__ nodes, int nInternal, float4* __restrict__ nodes4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nInternal) return;
    WChild c[4];
    int n = 2;
    wide_children(nodes, i, c[0], c[1]);
#pragma unroll
    for (int round = 0; round < 2; ++round) {
        int best = -1;
        float bestA = -1.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k < n && c[k].id >= 0) {
                const float a = wide_area(c[k]);
                if (a > bestA) {
                    bestA = a;
                    best = k;
                }
            }
        if (best < 0) break;
        WChild a, b;
        int id = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k == best) id = c[k].id;
        wide_children(nodes, id, a, b);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (k == best) c[k] = a;
            if (k == n) c[k] = b;
        }
        ++n;
    }
    float3 lo = c[0].lo, hi = c[0].hi;
#pragma unroll
    for (int k = 1; k < 4; ++k)
        if (k < n) {
            lo = fmin3(lo, c[k].lo);
            hi = fmax3(hi, c[k].hi);
        }
    const uint32_t ex = wide_exponent(lo.x, hi.x), ey = wide_exponent(lo.y, hi.y), ez = wide_exponent(lo.z, hi.z);
    const float sx = __uint_as_float(ex << 23), sy = __uint_as_float(ey << 23), sz = __uint_as_float(ez << 23);
    const float ix = __uint_as_float((254u - ex) << 23), iy = __uint_as_float((254u - ey) << 23), iz = __uint_as_float((254u - ez) << 23);   // 2^-(e)
    uint32_t q[6] = {0u, 0u, 0u, 0u, 0u, 0u};   // qlo.x, qlo.y, qlo.z, qhi.x, qhi.y, qhi.z: one byte per child
    int child[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        child[k] = kWideEmpty;
        if (k < n) {
            child[k] = c[k].id;
            q[0] |= wide_qlo(c[k].lo.x, lo.x, sx, ix) << (8 * k);
            q[1] |= wide_qlo(c[k].lo.y, lo.y, sy, iy) << (8 * k);
            q[2] |= wide_qlo(c[k].lo.z, lo.z, sz, iz) << (8 * k);
            q[3] |= wide_qhi(c[k].hi.x, lo.x, sx, ix) << (8 * k);
            q[4] |= wide_qhi(c[k].hi.y, lo.y, sy, iy) << (8 * k);
            q[5] |= wide_qhi(c[k].hi.z, lo.z, sz, iz) << (8 * k);
        }
    }
    float4* out = nodes4 + 4 * (size_t)i;
    out[0] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(ex | (ey << 8) | (ez << 16) | ((uint32_t)n << 24)));
    out[1] = make_float4(__int_as_float(child[0]), __int_as_float(child[1]), __int_as_float(child[2]), __int_as_float(child[3]));
    out[2] = make_float4(__uint_as_float(q[0]), __uint_as_float(q[1]), __uint_as_float(q[2]), __uint_as_float(q[3]));
    out[3] = make_float4(__uint_as_float(q[4]), __uint_as_float(q[5]), 0.0f, 0.0f);
}

// ---- export in the reference's 36-byte layout (include/bvh.cuh:34-58) --------------------------
struct RefNode36 {
    uint32_t parent;
    int32_t a, b;
    float mn[3], mx[3];
};
static_assert(sizeof(RefNode36) == 36, "reference BVHNode is 36 bytes");

__global__ void k_export_reference(const float4* __restrict__ nodes, const float4* __restrict__ triGeom, const int* __restrict__ parentInt,
                                   const int* __restrict__ parentLeaf, int nFaces, RefNode36* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int nInt = nFaces - 1;
    if (i < nInt) {
        const float4 n0 = nodes[4 * (size_t)i], n1 = nodes[4 * (size_t)i + 1], n2 = nodes[4 * (size_t)i + 2];
        const int4 ch = reinterpret_cast<const int4*>(nodes)[4 * (size_t)i + 3];
        RefNode36 r;
        r.parent = i == 0 ? 0u : (uint32_t)parentInt[i];
        r.a = ch.x >= 0 ? ch.x : (~ch.x) + nInt;
        r.b = ch.y >= 0 ? ch.y : (~ch.y) + nInt;
        r.mn[0] = fminf(n0.x, n1.x); r.mx[0] = fmaxf(n0.y, n1.y);
        r.mn[1] = fminf(n0.z, n1.z); r.mx[1] = fmaxf(n0.w, n1.w);
        r.mn[2] = fminf(n2.x, n2.z); r.mx[2] = fmaxf(n2.y, n2.w);
        out[i] = r;
    }
    if (i < nFaces) {
        RefNode36 r;
        const float4 c = triGeom[3 * (size_t)i + 2];
        r.a = __float_as_int(c.y);
        r.b = 0;
        if (nFaces > 1) {
            const int p = parentLeaf[i];
            r.parent = (uint32_t)p;
            const float4 n0 = nodes[4 * (size_t)p], n1 = nodes[4 * (size_t)p + 1], n2 = nodes[4 * (size_t)p + 2];
            const int4 ch = reinterpret_cast<const int4*>(nodes)[4 * (size_t)p + 3];
            const bool isLeft = ch.x == ~i;
            const float4 xy = isLeft ? n0 : n1;
            r.mn[0] = xy.x; r.mx[0] = xy.y; r.mn[1] = xy.z; r.mx[1] = xy.w;
            r.mn[2] = isLeft ? n2.x : n2.z; r.mx[2] = isLeft ? n2.y : n2.w;
        } else {
            const float4 a = triGeom[0], b = triGeom[1];
            const float3 v0 = f3(a.x, a.y, a.z), v1 = v0 + f3(a.w, b.x, b.y), v2 = v0 + f3(b.z, b.w, c.x);
            const float3 mn = fmin3(fmin3(v0, v1), v2), mx = fmax3(fmax3(v0, v1), v2);
            r.parent = 0;
            r.mn[0] = mn.x; r.mn[1] = mn.y; r.mn[2] = mn.z; r.mx[0] = mx.x; r.mx[1] = mx.y; r.mx[2] = mx.z;
        }
        out[nInt + i] = r;
    }
}

// ---- stats (only when the caller asks for them) --------------------------------------------------
__global__ void k_bvh_stats(const float4* __restrict__ nodes, const int* __restrict__ parentLeaf, const int* __restrict__ parentInt,
                            int nFaces, float* __restrict__ areaSum, uint32_t* __restrict__ maxDepth) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nFaces - 1) {
        const float4 n0 = nodes[4 * (size_t)i], n1 = nodes[4 * (size_t)i + 1], n2 = nodes[4 * (size_t)i + 2];
        const float ex = fmaxf(n0.y, n1.y) - fminf(n0.x, n1.x), ey = fmaxf(n0.w, n1.w) - fminf(n0.z, n1.z),
                    ez = fmaxf(n2.y, n2.w) - fminf(n2.x, n2.z);
        atomicAdd(areaSum + (i == 0 ? 0 : 1), 2.0f * (ex * ey + ey * ez + ez * ex));
        // children areas that are leaves
        const float lx = n0.y - n0.x, ly = n0.w - n0.z, lz = n2.y - n2.x, rx = n1.y - n1.x, ry = n1.w - n1.z, rz = n2.w - n2.z;
        const int4 ch = reinterpret_cast<const int4*>(nodes)[4 * (size_t)i + 3];
        float leafArea = 0.f;
        if (ch.x < 0) leafArea += 2.0f * (lx * ly + ly * lz + lz * lx);
        if (ch.y < 0) leafArea += 2.0f * (rx * ry + ry * rz + rz * rx);
        if (leafArea > 0.f) atomicAdd(areaSum + 2, leafArea);
    }
    if (i < nFaces && nFaces > 1) {
        uint32_t depth = 1;
        int cur = parentLeaf[i];
        while (cur > 0) {
            cur = parentInt[cur];
            ++depth;
        }
        atomicMax(maxDepth, depth);
    }
}

}   // namespace tpt

using namespace tpt;

constexpr size_t kLazyNormalFaces = (size_t)1 << 16;
static std::atomic<uint64_t> g_geom_epoch{0};   // one counter for all scenes and contexts
static inline unsigned grid_for(size_t n, int block) { return (unsigned)((n + block - 1) / block); }

int tpt_scene_build_tables(tpt_scene* s) {
    cudaStream_t st = s->ctx->stream;
    TPT_CUDA(cudaMemsetAsync(s->vert_obj, 0xff, s->n_vertices * sizeof(uint32_t), st));
    k_face_tables<<<grid_for(s->n_faces, 256), 256, 0, st>>>(s->indices, s->objects, (int)s->n_objects, (uint32_t)s->n_faces, s->vert_obj,
                                                             s->face_mtl);
    TPT_CUDA(cudaGetLastError());
    s->ctx->kernel_launches += 1;
    return TPT_OK;
}

int tpt_bvh_build_impl(tpt_scene* s) {
    tpt_context* ctx = s->ctx;
    cudaStream_t st = ctx->stream;
    const int F = (int)s->n_faces;
    const uint32_t V = (uint32_t)s->n_vertices;
    const int persistent = kNumSMs * 8;
    const bool lazyNormals = s->n_faces >= kLazyNormalFaces;   // large scenes: normal records on first render (tpt_scene_ensure_normals)

    k_init_bounds<<<1, 32, 0, st>>>(s->bounds);
    (lazyNormals ? k_transform<false> : k_transform<true>)<<<min(persistent, (int)grid_for(V, 256)), 256, 0, st>>>(
        s->obj_verts, s->obj_norms, s->vert_obj, s->vert_trans, s->normal_trans, V, s->world_verts, s->world_norms, s->bounds);
    // 63-bit keys like the reference (21 bits per axis) below 2^20 triangles.  From there on the radix passes are what the keys cost, and the
    // tree does not need the resolution: 10 bits per axis = 30-bit keys = 4 passes instead of 8; 2^30 cells are >= 64 per triangle up to
    // 2^24 triangles, equal keys are ordered by index.  Measured on the 10 M-triangle soup with the search-based hierarchy (bits per axis:
    // build ms | primary, secondary Mrays/s | SAH cost): 13: 1.905 | 2 269, 1 275 | 149.1; 11 (33 bits cut to 32): 1.725 | 2 257, 1 277 | 149.2;
    // 10: 1.697 | 2 258, 1 280 | 149.4; 9: 1.650 | 2 226, 1 238 | 152.5; 8 (3 passes): 1.477 | 2 162, 1 204 | 159.9 — 10 is the narrowest key
    // that costs the rays nothing.  Above 2^24 triangles: 13 bits, 5 passes.
    static const int bitsEnv = getenv("TPT_MORTON_BITS_LARGE") ? atoi(getenv("TPT_MORTON_BITS_LARGE")) : 0;   // tuning switch (8..21)
    const int bitsPerAxis = F < (1 << 20) ? 21 : bitsEnv > 0 ? bitsEnv : F <= (1 << 24) ? 10 : 13;
    const int keyBits = 3 * bitsPerAxis - morton_key_shift(bitsPerAxis);
    // Sorted (key, face) pairs must end in keys[1] / vals[1]; keys[0] / vals[0] are scratch afterwards.
    const int nPasses = (keyBits + 7) / 8;   // 4, 5 or 8
    const int first = (nPasses & 1) ? 0 : 1; // the ping-pong starts in the buffer that makes it END in buffer 1
    if (F <= kSortTile) {
        k_sort_small<<<1, kSortThreads, 0, st>>>(s->indices, s->world_verts, s->bounds, (uint32_t)F, bitsPerAxis, nPasses, s->keys[1], s->vals[1]);
    } else {
        const uint32_t nTiles = (uint32_t)((F + kSortTile - 1) / kSortTile);
        const size_t histBytes = (size_t)kSortMaxPasses * 256 * sizeof(uint32_t), ticketBytes = 64;
        const size_t need = histBytes + ticketBytes + (size_t)nPasses * nTiles * 256 * sizeof(uint32_t);
        if (need > s->sort_tmp_bytes) {
            TPT_CUDA(cudaStreamSynchronize(st));
            if (s->sort_tmp) cudaFree(s->sort_tmp);
            s->sort_tmp = nullptr;
            s->sort_tmp_bytes = 0;
            TPT_CUDA(cudaMalloc(&s->sort_tmp, need));
            s->sort_tmp_bytes = need;
        }
        uint32_t* hist = static_cast<uint32_t*>(s->sort_tmp);
        uint32_t* ticket = hist + kSortMaxPasses * 256;
        uint32_t* status = ticket + ticketBytes / sizeof(uint32_t);
        TPT_CUDA(cudaMemsetAsync(s->sort_tmp, 0, need, st));
        k_morton_hist<<<min(persistent, (int)grid_for(F, 256)), 256, 0, st>>>(s->indices, s->world_verts, s->bounds, (uint32_t)F, bitsPerAxis, nPasses,
                                                                                s->keys[first], s->vals[first], hist);
        const unsigned grid = (unsigned)std::min<uint32_t>(nTiles, (uint32_t)(kNumSMs * 4));
        for (int p = 0; p < nPasses; ++p) {
            const int in = first ^ (p & 1), out = in ^ 1;
            k_onesweep<<<grid, kSortThreads, 0, st>>>(s->keys[in], s->vals[in], s->keys[out], s->vals[out], (uint32_t)F, 8 * p, hist + p * 256,
                                                      status + (size_t)p * nTiles * 256, ticket + p, nTiles);
        }
    }
    TPT_CUDA(cudaGetLastError());
    // vals[0] / keys[0] were the sort's input and are free now: range ends and the walk's frontier live there
    int* frontier = reinterpret_cast<int*>(s->keys[0]);
    int* rangeOther = reinterpret_cast<int*>(s->vals[0]);
    if (F > 1) TPT_CUDA(cudaMemsetAsync(s->flags, 0, (size_t)(F - 1) * sizeof(uint32_t), st));
    static const bool karras = getenv("TPT_BUILD_KARRAS") != nullptr;   // test / tuning switch: the search-based hierarchy + refit walk
    if (karras) {
        if (F > 1) k_hierarchy<<<grid_for(F - 1, 256), 256, 0, st>>>(s->keys[1], F, s->nodes, s->parent_int, s->parent_leaf, rangeOther);
        k_leaves_refit<<<grid_for(F, kRefitBlock), kRefitBlock, 0, st>>>(s->vals[1], s->indices, s->world_verts, s->world_norms, s->face_mtl, F,
                                                                         s->parent_int, s->parent_leaf, rangeOther,
                                                                         s->flags, s->nodes, s->tri_geom, lazyNormals ? nullptr : s->tri_norm, frontier, s->bounds + 6);
        if (F > 1) k_refit_upper<<<grid_for((size_t)F / 2 + 1, 256), 256, 0, st>>>(frontier, s->bounds + 6, s->parent_int, s->flags, s->nodes);
    } else {
        k_leaves_build<<<grid_for(F, kRefitBlock), kRefitBlock, 0, st>>>(s->keys[1], s->vals[1], s->indices, s->world_verts, s->world_norms, s->face_mtl, F,
                                                                         s->parent_int, s->parent_leaf, rangeOther, s->flags, s->merge_slots, s->nodes,
                                                                         s->tri_geom, lazyNormals ? nullptr : s->tri_norm, frontier, s->bounds + 6);
        if (F > 1) k_build_upper<<<(unsigned)std::min<size_t>(grid_for((size_t)F / 16 + 1, 256), (size_t)kNumSMs * 8), 256, 0, st>>>(s->keys[1], F, frontier, s->bounds + 6, s->parent_int, s->parent_leaf,
                                                                                  rangeOther, s->flags, s->merge_slots, s->nodes);
    }
    if (F > 1 && s->nodes4) {
        k_collapse4<<<grid_for((size_t)F - 1, 256), 256, 0, st>>>(s->nodes, F - 1, s->nodes4);
        ctx->kernel_launches += 1;
    }
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += (karras ? 5 : 4) + (F <= kSortTile ? 1 : 1 + nPasses);
    s->built = true;
    s->geom_epoch = ++g_geom_epoch;   // unique across scenes: a new scene at a recycled address never matches a cached key
    if (!lazyNormals) s->norm_epoch = s->geom_epoch;
    s->topology = true;
    return TPT_OK;
}

// Refit: new object->world matrices (tpt_scene_set_transforms) on an unchanged mesh.  The reference rebuilds everything per
// frame (src/path_tracer.cu:540-542); here the sorted leaf order, children, parents and range ends of the last full build are
// kept and only the geometry-dependent part runs again: transform, triangle/normal records, bottom-up boxes.  No Morton keys,
// no sort, no hierarchy.  The boxes are again the exact bounds of their triangles, so traversal results are those of a rebuild
// (closest hit under the tie rule does not depend on the tree shape); only the tree quality can drift with large motion.
int tpt_bvh_refit_impl(tpt_scene* s) {
    tpt_context* ctx = s->ctx;
    cudaStream_t st = ctx->stream;
    const int F = (int)s->n_faces;
    const uint32_t V = (uint32_t)s->n_vertices;
    const int persistent = kNumSMs * 8;
    const bool lazyNormals = s->n_faces >= kLazyNormalFaces;   // large scenes: normal records on first render (tpt_scene_ensure_normals)
    k_init_bounds<<<1, 32, 0, st>>>(s->bounds);
    (lazyNormals ? k_transform<false> : k_transform<true>)<<<min(persistent, (int)grid_for(V, 256)), 256, 0, st>>>(
        s->obj_verts, s->obj_norms, s->vert_obj, s->vert_trans, s->normal_trans, V, s->world_verts, s->world_norms, s->bounds);
    if (F > 1) TPT_CUDA(cudaMemsetAsync(s->flags, 0, (size_t)(F - 1) * sizeof(uint32_t), st));
    int* frontier = reinterpret_cast<int*>(s->keys[0]);
    k_leaves_refit<<<grid_for(F, kRefitBlock), kRefitBlock, 0, st>>>(s->vals[1], s->indices, s->world_verts, s->world_norms, s->face_mtl, F,
                                                                     s->parent_int, s->parent_leaf, reinterpret_cast<const int*>(s->vals[0]),
                                                                     s->flags, s->nodes, s->tri_geom, lazyNormals ? nullptr : s->tri_norm, frontier, s->bounds + 6);
    if (F > 1) k_refit_upper<<<grid_for((size_t)F / 2 + 1, 256), 256, 0, st>>>(frontier, s->bounds + 6, s->parent_int, s->flags, s->nodes);
    if (F > 1 && s->nodes4) {
        k_collapse4<<<grid_for((size_t)F - 1, 256), 256, 0, st>>>(s->nodes, F - 1, s->nodes4);
        ctx->kernel_launches += 1;
    }
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += 4;
    s->built = true;
    s->geom_epoch = ++g_geom_epoch;   // unique across scenes: a new scene at a recycled address never matches a cached key
    if (!lazyNormals) s->norm_epoch = s->geom_epoch;
    return TPT_OK;
}

int tpt_scene_ensure_normals(tpt_scene* s) {
    if (s->norm_epoch == s->geom_epoch) return TPT_OK;
    const int F = (int)s->n_faces;
    const uint32_t V = (uint32_t)s->n_vertices;
    k_transform_normals<<<min(kNumSMs * 8, (int)grid_for(V, 256)), 256, 0, s->ctx->stream>>>(s->obj_norms, s->vert_obj, s->normal_trans, V, s->world_norms);
    k_leaf_normals<<<grid_for(F, 256), 256, 0, s->ctx->stream>>>(s->vals[1], s->indices, s->world_norms, F, s->tri_norm);
    TPT_CUDA(cudaGetLastError());
    s->ctx->kernel_launches += 2;
    s->norm_epoch = s->geom_epoch;
    return TPT_OK;
}

int tpt_bvh_stats_impl(tpt_scene* s, tpt_bvh_stats* stats) {
    cudaStream_t st = s->ctx->stream;
    const int F = (int)s->n_faces;
    float* d_area = nullptr;
    TPT_CUDA(cudaMalloc(&d_area, 4 * sizeof(float)));
    TPT_CUDA(cudaMemsetAsync(d_area, 0, 4 * sizeof(float), st));
    uint32_t* d_depth = reinterpret_cast<uint32_t*>(d_area + 3);
    if (F > 1) k_bvh_stats<<<grid_for(F, 256), 256, 0, st>>>(s->nodes, s->parent_leaf, s->parent_int, F, d_area, d_depth);
    float h[4];
    uint32_t hb[6];
    TPT_CUDA(cudaMemcpyAsync(h, d_area, sizeof h, cudaMemcpyDeviceToHost, st));
    TPT_CUDA(cudaMemcpyAsync(hb, s->bounds, sizeof hb, cudaMemcpyDeviceToHost, st));
    TPT_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_area);
    stats->n_faces = (uint32_t)F;
    stats->n_nodes = (uint32_t)(F > 0 ? F - 1 : 0);
    memcpy(&stats->max_depth, &h[3], 4);
    // SAH with Ct = 1.2, Ci = 1: (sum internal areas * Ct + sum leaf areas * Ci) / root area
    stats->sah_cost = h[0] > 0.f ? (1.2f * (h[0] + h[1]) + h[2]) / h[0] : 0.f;
    for (int k = 0; k < 3; ++k) {
        stats->bounds_min[k] = ord2f(hb[k]);
        stats->bounds_max[k] = ord2f(hb[3 + k]);
    }
    return TPT_OK;
}

int tpt_bvh_export_reference_impl(tpt_scene* s, void* out, size_t nNodes) {
    cudaStream_t st = s->ctx->stream;
    const int F = (int)s->n_faces;
    RefNode36* d = nullptr;
    TPT_CUDA(cudaMalloc(&d, nNodes * sizeof(RefNode36)));
    k_export_reference<<<grid_for(F, 256), 256, 0, st>>>(s->nodes, s->tri_geom, s->parent_int, s->parent_leaf, F, d);
    cudaError_t e = cudaMemcpyAsync(out, d, nNodes * sizeof(RefNode36), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d);
    TPT_CUDA(e);
    return TPT_OK;
}
