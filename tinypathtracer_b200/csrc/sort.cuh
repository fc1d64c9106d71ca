// sort.cuh — radix sort of the (Morton key64, face32) pairs of the LBVH build, written for sm_100a.
//
// Replaces thrust::sort_by_key of the reference (src/bvh.cu:326, which moves whole 36-byte nodes through every radix pass) and
// the cub::DeviceRadixSort call this file's first version of the build used.  Least-significant-digit radix sort, 8-bit digits,
// every pass STABLE, so the result is the unique stable order: equal keys keep ascending face index.
//
//   small inputs (F <= kSortTile): ONE kernel, one block — Morton keys are produced in registers, all passes run in shared
//       memory, only the sorted pairs are written.  Every bundled scene takes this path (1 launch instead of key kernel + sort).
//   large inputs: "onesweep" (Adinets & Merrill 2022):
//       k_morton_hist   produces the keys where they are first needed, writes them once, and builds the digit histograms of ALL
//                       passes in the same sweep (shared-memory counters, one global atomic per bin and block);
//       k_onesweep      one launch per pass; a block takes tiles of kSortTile pairs in arrival order (atomic ticket), ranks
//                       them by digit with warp match operations, publishes its per-digit counts and obtains its global
//                       offsets by decoupled look-back over the tiles before it (no separate scan pass, each pair is read
//                       once and written once per pass), reorders the tile in shared memory and writes runs of equal digits.
//       Traffic per pass: 12 B read + 12 B written per pair — the figure SURVEY §8(d) counts.
#pragma once
#include "scene.cuh"

namespace tpt {

// ---- ordered-uint encoding of floats for atomicMin/Max ---------------------------------------
__host__ __device__ __forceinline__ uint32_t f2ord(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

#ifndef TPT_SORT_MIN_BLOCKS
#define TPT_SORT_MIN_BLOCKS 3
#endif
constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;                            // pairs per thread
constexpr int kSortTile = kSortThreads * kSortItems;      // 4096 pairs: 32 KB of keys in shared memory
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortWarpItems = 32 * kSortItems;           // a warp owns 512 consecutive pairs of the tile
constexpr int kSortMaxPasses = 8;
constexpr uint32_t kFlagAgg = 1u << 30, kFlagPrefix = 2u << 30, kFlagMask = 3u << 30;

struct SortSmem {
    unsigned long long keys[kSortTile];                   // exchange buffer (keys, then reused for the values)
    uint32_t warpHist[kSortWarps][256];                   // per-warp digit counters -> exclusive offsets of the warp inside the tile
    uint32_t digitStart[256];                             // tile-local exclusive prefix over digits
    uint32_t globalBase[256];                             // where the tile's run of each digit starts in the output
    uint32_t ticket;
};

// Stable ranks of one warp's 16 x 32 pairs by digit.  Item k of lane l is pair (k * 32 + l) of the warp's segment; ranks follow that
// order.  wh = this warp's 256 counters (zeroed).  Invalid pairs (beyond the input) take no rank.
__device__ __forceinline__ void warp_rank(const unsigned long long (&key)[kSortItems], const bool (&valid)[kSortItems], int shift, uint32_t* wh,
                                          uint32_t (&rank)[kSortItems]) {
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t d = valid[k] ? (uint32_t)(key[k] >> shift) & 255u : 256u;
        const unsigned m = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(m) - 1;
        uint32_t base = 0;
        if ((int)lane == leader && d < 256u) {
            base = wh[d];
            wh[d] = base + (uint32_t)__popc(m);
        }
        base = __shfl_sync(0xffffffffu, base, leader);
        rank[k] = base + (uint32_t)__popc(m & lt);
        __syncwarp();
    }
}

// After warp_rank of every warp: turn warpHist[w][d] (counts) into exclusive offsets across the warps, and return digit d's
// total in the tile to thread d.  Needs __syncthreads() before and after.
__device__ __forceinline__ uint32_t warp_offsets(SortSmem& sm) {
    const uint32_t d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
        const uint32_t c = sm.warpHist[w][d];
        sm.warpHist[w][d] = run;
        run += c;
    }
    return run;
}

// Exclusive prefix sum over the 256 digits (thread d holds `count`), result in sm.digitStart[d].  All 256 threads.
__device__ __forceinline__ void digit_scan(SortSmem& sm, uint32_t count) {
    __shared__ uint32_t warpSum[kSortWarps];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (uint32_t)o) inc += v;
    }
    if (lane == 31) warpSum[w] = inc;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int k = 0; k < kSortWarps; ++k)
        if ((uint32_t)k < w) before += warpSum[k];
    sm.digitStart[threadIdx.x] = before + inc - count;
    __syncthreads();
}

// Morton key of face f (k_morton's arithmetic; bvh.cu keeps the scalar kernel's description).
struct MortonParams {
    float3 bmin, scale;
    float qmax;
    int shift;   // low bits of the interleaved key that are dropped (x's, then y's last bit) so that the key fills whole 8-bit digits
};
// 3 * bits that is 1 or 2 above a multiple of 8 (11 bits per axis: 33) is cut down to the multiple: one radix pass fewer for the last bit of one axis
__host__ __device__ __forceinline__ int morton_key_shift(int bitsPerAxis) { return ((3 * bitsPerAxis) & 7) <= 2 ? ((3 * bitsPerAxis) & 7) : 0; }
__device__ __forceinline__ uint64_t sort_expand21(uint64_t x) {
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
__device__ __forceinline__ unsigned long long face_key(const uint32_t* __restrict__ indices, const float* __restrict__ worldVerts, const MortonParams& mp,
                                                       uint32_t f) {
    const uint32_t i0 = indices[3 * f], i1 = indices[3 * f + 1], i2 = indices[3 * f + 2];
    const float3 a = f3(worldVerts[3 * i0], worldVerts[3 * i0 + 1], worldVerts[3 * i0 + 2]);
    const float3 b = f3(worldVerts[3 * i1], worldVerts[3 * i1 + 1], worldVerts[3 * i1 + 2]);
    const float3 c = f3(worldVerts[3 * i2], worldVerts[3 * i2 + 1], worldVerts[3 * i2 + 2]);
    const float3 ctr = (fmin3(fmin3(a, b), c) + fmax3(fmax3(a, b), c)) * 0.5f;   // AABB centre, as reference bvh.cu:139-142
    const float3 q = (ctr - mp.bmin) * mp.scale;
    const uint64_t qx = (uint64_t)fminf(fmaxf(q.x, 0.f), mp.qmax);
    const uint64_t qy = (uint64_t)fminf(fmaxf(q.y, 0.f), mp.qmax);
    const uint64_t qz = (uint64_t)fminf(fmaxf(q.z, 0.f), mp.qmax);
    return (sort_expand21(qx) | (sort_expand21(qy) << 1) | (sort_expand21(qz) << 2)) >> mp.shift;
}
__device__ __forceinline__ MortonParams morton_params(const uint32_t* __restrict__ bounds, int bitsPerAxis) {
    const float3 bmin = f3(ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2]));
    const float3 bmax = f3(ord2f(bounds[3]), ord2f(bounds[4]), ord2f(bounds[5]));
    const float3 ext = bmax - bmin;
    const float k = (float)(1u << bitsPerAxis);   // cells per axis (2^21: the reference's key width; 2^16 for huge scenes)
    MortonParams mp;
    mp.bmin = bmin;
    mp.scale = f3(ext.x > 0.f ? k / ext.x : 0.f, ext.y > 0.f ? k / ext.y : 0.f, ext.z > 0.f ? k / ext.z : 0.f);
    mp.qmax = k - 1.0f;
    mp.shift = morton_key_shift(bitsPerAxis);
    return mp;
}

// ---- small inputs: everything in one block --------------------------------------------------------------------------
// The pairs are dealt evenly: with `items` = ceil(F / 256) pairs per thread, warp w owns pairs [w * 32 * items, (w + 1) * 32 * items)
// and item k of lane l is pair w * 32 * items + k * 32 + l, so all eight warps rank (ball.gltf: 5 rounds of match instead of 16 in
// three warps).  Rounds k >= items hold no pair and are skipped by the whole block.
__device__ __forceinline__ void warp_rank_n(const unsigned long long (&key)[kSortItems], const bool (&valid)[kSortItems], int items, int shift,
                                            uint32_t* wh, uint32_t (&rank)[kSortItems]) {
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        if (k < items) {   // uniform
            const uint32_t d = valid[k] ? (uint32_t)(key[k] >> shift) & 255u : 256u;
            const unsigned m = __match_any_sync(0xffffffffu, d);
            const int leader = __ffs(m) - 1;
            uint32_t base = 0;
            if ((int)lane == leader && d < 256u) {
                base = wh[d];
                wh[d] = base + (uint32_t)__popc(m);
            }
            base = __shfl_sync(0xffffffffu, base, leader);
            rank[k] = base + (uint32_t)__popc(m & lt);
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(kSortThreads) k_sort_small(const uint32_t* __restrict__ indices, const float* __restrict__ worldVerts,
                                                             const uint32_t* __restrict__ bounds, uint32_t nFaces, int bitsPerAxis, int nPasses,
                                                             unsigned long long* __restrict__ keysOut, uint32_t* __restrict__ valsOut) {
    __shared__ SortSmem sm;
    uint32_t* sv = reinterpret_cast<uint32_t*>(sm.keys);   // the values take the same route through the same buffer, after the keys
    const MortonParams mp = morton_params(bounds, bitsPerAxis);
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int items = (int)((nFaces + kSortThreads - 1) / kSortThreads);
    const uint32_t warpBase = w * 32u * (uint32_t)items;
    unsigned long long key[kSortItems];
    uint32_t val[kSortItems], rank[kSortItems];
    bool valid[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t f = warpBase + k * 32 + lane;
        valid[k] = k < items && f < nFaces;
        val[k] = f;
        key[k] = valid[k] ? face_key(indices, worldVerts, mp, f) : ~0ull;
        rank[k] = 0;
    }
    for (int p = 0; p < nPasses; ++p) {
        for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&sm.warpHist[0][0])[i] = 0u;
        __syncthreads();
        warp_rank_n(key, valid, items, 8 * p, sm.warpHist[w], rank);
        __syncthreads();
        const uint32_t count = warp_offsets(sm);
        __syncthreads();
        digit_scan(sm, count);
#pragma unroll
        for (int k = 0; k < kSortItems; ++k)
            if (valid[k]) {
                const uint32_t d = (uint32_t)(key[k] >> (8 * p)) & 255u;
                rank[k] = sm.digitStart[d] + sm.warpHist[w][d] + rank[k];   // position in the reordered tile
                sm.keys[rank[k]] = key[k];
            }
        __syncthreads();
        // back to the thread arrangement (valid[] is the same before and after: it only depends on the pair's index)
#pragma unroll
        for (int k = 0; k < kSortItems; ++k)
            if (valid[k]) key[k] = sm.keys[warpBase + k * 32 + lane];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kSortItems; ++k)
            if (valid[k]) sv[rank[k]] = val[k];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kSortItems; ++k)
            if (valid[k]) val[k] = sv[warpBase + k * 32 + lane];
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kSortItems; ++k)
        if (valid[k]) {
            keysOut[warpBase + k * 32 + lane] = key[k];
            valsOut[warpBase + k * 32 + lane] = val[k];
        }
}

// ---- large inputs ------------------------------------------------------------------------------------------------------
// Keys + identity values + the digit histograms of all passes in one sweep over the faces.
__global__ void __launch_bounds__(256) k_morton_hist(const uint32_t* __restrict__ indices, const float* __restrict__ worldVerts,
                                                     const uint32_t* __restrict__ bounds, uint32_t nFaces, int bitsPerAxis, int nPasses,
                                                     unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ hist) {
    __shared__ uint32_t sHist[kSortMaxPasses][256];
    for (int i = threadIdx.x; i < kSortMaxPasses * 256; i += blockDim.x) (&sHist[0][0])[i] = 0u;
    __syncthreads();
    const MortonParams mp = morton_params(bounds, bitsPerAxis);
    for (uint32_t f = blockIdx.x * blockDim.x + threadIdx.x; f < nFaces; f += gridDim.x * blockDim.x) {
        const unsigned long long key = face_key(indices, worldVerts, mp, f);
        keys[f] = key;
        vals[f] = f;
        for (int p = 0; p < nPasses; ++p) atomicAdd(&sHist[p][(uint32_t)(key >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nPasses * 256; i += blockDim.x) {
        const uint32_t c = (&sHist[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
}

// One pass.  hist = this pass's 256 global digit counts; status = nTiles x 256 words (zeroed), tile t digit d at [t * 256 + d]:
// count | kFlagAgg once the tile's own count is known, inclusive prefix | kFlagPrefix once everything before it is too.
__global__ void __launch_bounds__(kSortThreads, TPT_SORT_MIN_BLOCKS) k_onesweep(const unsigned long long* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                           unsigned long long* __restrict__ keysOut, uint32_t* __restrict__ valsOut, uint32_t n, int shift,
                                                           const uint32_t* __restrict__ hist, volatile uint32_t* status, uint32_t* ticket, uint32_t nTiles) {
    __shared__ SortSmem sm;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, d = threadIdx.x;
    // exclusive scan of the global digit counts: where each digit's run starts in the output
    digit_scan(sm, hist[d]);
    const uint32_t digitBase = sm.digitStart[d];
    __syncthreads();
    while (true) {
        if (threadIdx.x == 0) sm.ticket = atomicAdd(ticket, 1u);
        for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&sm.warpHist[0][0])[i] = 0u;
        __syncthreads();
        const uint32_t tile = sm.ticket;
        if (tile >= nTiles) break;
        const uint32_t tileBase = tile * (uint32_t)kSortTile;
        unsigned long long key[kSortItems];
        uint32_t rank[kSortItems];
        bool valid[kSortItems];
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t j = tileBase + w * kSortWarpItems + k * 32 + lane;
            valid[k] = j < n;
            key[k] = valid[k] ? __ldcs(keysIn + j) : ~0ull;
        }
        warp_rank(key, valid, shift, sm.warpHist[w], rank);
        __syncthreads();
        const uint32_t count = warp_offsets(sm);
        // publish this tile's count of digit d, then look back for what lies before the tile
        volatile uint32_t* mine = status + (size_t)tile * 256u + d;
        uint32_t excl = 0;
        if (tile == 0) {
            *mine = count | kFlagPrefix;
        } else {
            *mine = count | kFlagAgg;
            for (int t = (int)tile - 1; t >= 0; --t) {
                uint32_t v;
                do {
                    v = status[(size_t)t * 256u + d];
                } while ((v & kFlagMask) == 0u);
                excl += v & ~kFlagMask;
                if (v & kFlagPrefix) break;
            }
            *mine = (excl + count) | kFlagPrefix;
        }
        sm.globalBase[d] = digitBase + excl;
        __syncthreads();          // warpHist offsets + globalBase visible
        digit_scan(sm, count);    // tile-local starts of the digits
        // reorder the keys in shared memory, then write them out in runs of equal digits
#pragma unroll
        for (int k = 0; k < kSortItems; ++k)
            if (valid[k]) {
                const uint32_t dk = (uint32_t)(key[k] >> shift) & 255u;
                rank[k] = sm.digitStart[dk] + sm.warpHist[w][dk] + rank[k];   // position inside the reordered tile
                sm.keys[rank[k]] = key[k];
            }
        __syncthreads();
        const uint32_t tileCount = min((uint32_t)kSortTile, n - tileBase);
        uint32_t dst[kSortItems];
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t j = k * kSortThreads + threadIdx.x;
            if (j < tileCount) {
                const unsigned long long kk = sm.keys[j];
                const uint32_t dk = (uint32_t)(kk >> shift) & 255u;
                dst[k] = sm.globalBase[dk] + (j - sm.digitStart[dk]);
                keysOut[dst[k]] = kk;
            }
        }
        __syncthreads();
        // the values take the same route through the same buffer
        uint32_t* sv = reinterpret_cast<uint32_t*>(sm.keys);
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t j = tileBase + w * kSortWarpItems + k * 32 + lane;
            if (valid[k]) sv[rank[k]] = __ldcs(valsIn + j);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t j = k * kSortThreads + threadIdx.x;
            if (j < tileCount) valsOut[dst[k]] = sv[j];
        }
        __syncthreads();
    }
}

}   // namespace tpt
