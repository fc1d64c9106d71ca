// comm.cu — the one exchange step of the multi-GPU path, for a host that drives all GPUs from ONE process (as the
// reference drives its single GPU: one host thread, src/path_tracer.cu:491-554): the sum of the per-GPU accumulation
// buffers.  No library collective: every GPU reduces its own 1/n slice by reading the peers' buffers
// directly over NVLink (peer loads in k_reduce_slice), then collects the other reduced slices with peer copies.
// The buffers are int64 fixed point, so the sum is exact and the result does not depend on n or on the order.
// One process PER GPU (bench.py under torchrun): the frame exchange at the end of this file — the same reduction FUSED with the
// resolve, over CUDA IPC mappings of the peers' buffers: k_xchg_resolve sums this rank's 1/n of the pixels straight out of the n
// accumulators, converts them to BGRA8 and stores them into the root rank's framebuffer, so 24 bytes per pixel cross NVLink
// once (reduce-scatter) and 4 bytes come back, instead of the 2 x 24 of an all-reduce followed by n full resolves.  (A host
// that prefers a library collective wraps the accumulator in a tensor instead: tpt_accum_wrap / tpt_accum_device_ptr.)
#include <cuda.h>
#include <unistd.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "scene.cuh"

namespace tpt {

constexpr int kMaxPeers = 16;
struct PeerPtrs {
    const unsigned long long* p[kMaxPeers];
};

// out[i] = sum over the n buffers of p[k][i], i in [begin, end); out is this GPU's own buffer (one of the p[k]): only
// this GPU reads or writes that slice of it during the launch (it aliases src.p[r]: no __restrict__).  128-bit loads — every
// buffer is 16-byte aligned (cudaMalloc, or checked by tpt_accum_wrap) —, the peers' data arrives over NVLink.
__global__ void __launch_bounds__(256) k_reduce_slice(PeerPtrs src, int n, unsigned long long* out, size_t begin, size_t end) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t b2 = (begin + 1) & ~(size_t)1, e2 = end & ~(size_t)1;   // 16-byte aligned middle part
    for (size_t i = b2 / 2 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < e2 / 2; i += stride) {
        ulonglong2 acc = make_ulonglong2(0ull, 0ull);
        for (int k = 0; k < n; ++k) {
            const ulonglong2 v = reinterpret_cast<const ulonglong2*>(src.p[k])[i];
            acc.x += v.x;
            acc.y += v.y;
        }
        reinterpret_cast<ulonglong2*>(out)[i] = acc;
    }
    if (blockIdx.x == 0 && threadIdx.x < 2) {   // at most one unaligned element at each end
        const size_t i = threadIdx.x == 0 ? begin : end - 1;
        const bool edge = threadIdx.x == 0 ? (begin < b2 && begin < end) : (e2 < end && end - 1 >= b2 && end > begin);
        if (edge) {
            unsigned long long acc = 0ull;
            for (int k = 0; k < n; ++k) acc += src.p[k][i];
            out[i] = acc;
        }
    }
}


// ---- frame exchange between processes (one per GPU) ---------------------------------------------------------------------
// Every rank owns one "shared block": XchgFlags, then its framebuffer.  Peers map it (and the accumulator) through CUDA IPC.
// Flags are sequence numbers (frame s = 1, 2, ... of this exchange), written by the peer named by the index with a
// system-scope release store and polled by the owner with system-scope acquire loads; a value only ever grows.
struct XchgFlags {
    unsigned int ready[kMaxPeers];   // ready[p] = s: rank p's samples of frame s are complete in its accumulator
    unsigned int done[kMaxPeers];    // done[p]  = s: rank p has read its pixels of frame s from every accumulator and stored them
    unsigned int blocks_done;        // owner only: completion count of k_xchg_resolve
    unsigned int error;              // owner only: a wait gave up (kXchgTimeout)
};
constexpr size_t kXchgFbOffset = 256;
static_assert(sizeof(XchgFlags) <= kXchgFbOffset, "flags, then the framebuffer at a 256-byte boundary");
struct XchgPeers {
    XchgFlags* flags[kMaxPeers];
    const unsigned long long* acc[kMaxPeers];
    uchar4* fb[kMaxPeers];
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// poll *flag until it reaches seq (wrap-safe); false after timeoutNs: the caller records the error and goes on, so that a
// rank that died cannot hang the others' GPUs
__device__ __forceinline__ bool xchg_wait(const unsigned int* flag, unsigned int seq, unsigned long long timeoutNs) {
    const unsigned long long t0 = globaltimer_ns();
    while ((int)(ld_acquire_sys(flag) - seq) < 0) {
        if (globaltimer_ns() - t0 > timeoutNs) return false;
        __nanosleep(200);
    }
    return true;
}

// thread p: tell rank p that this rank's frame `seq` is in its accumulator (everything queued before this launch on the
// stream has completed), then wait until rank p says the same.  One block of kMaxPeers threads.
__global__ void k_xchg_ready(XchgPeers peers, int rank, int world, unsigned int seq, unsigned long long timeoutNs) {
    const int p = threadIdx.x;
    if (p >= world) return;
    __threadfence_system();
    st_release_sys(&peers.flags[p]->ready[rank], seq);
    if (!xchg_wait(&peers.flags[rank]->ready[p], seq, timeoutNs)) peers.flags[rank]->error = 1u;
}

// thread p: wait until rank p has finished its part of frame `seq` — after this the owner may clear its accumulator (nobody reads it
// any more) and, on the root, the framebuffer is complete.
__global__ void k_xchg_wait_done(XchgPeers peers, int rank, int world, unsigned int seq, unsigned long long timeoutNs) {
    const int p = threadIdx.x;
    if (p >= world) return;
    if (!xchg_wait(&peers.flags[rank]->done[p], seq, timeoutNs)) peers.flags[rank]->error = 1u;
}

// Pixels [begin, end) of the frame (accumulator order: row 0 = bottom): sum of the n accumulators -> copyToFB's bytes (k_resolve,
// render.cu) -> framebuffer of rank `root`, or of every rank if root < 0.  A block handles kXchgTile pixels at a time: the
// 3 x int64 per pixel are fetched as a flat run of 128-bit words (coalesced: 512 contiguous bytes per warp and peer, all peers'
// loads of a word in flight together), summed into shared memory, then resolved two pixels per thread.  begin and the tile
// size are even, so every tile starts on a 16-byte boundary.  The last block to finish publishes done[rank] = seq to all ranks.
constexpr int kXchgTile = 512;   // pixels per block and pass = 768 x 16 bytes
__global__ void __launch_bounds__(256) k_xchg_resolve(XchgPeers peers, int rank, int world, int root, unsigned int seq, uint32_t begin, uint32_t end,
                                                      int W, int H, float invSpp, unsigned int alpha) {
    __shared__ ulonglong2 sSum[kXchgTile * 3 / 2];
    __shared__ bool sLast;
    const unsigned long long* sWords = reinterpret_cast<const unsigned long long*>(sSum);
    for (uint32_t tile = begin + blockIdx.x * kXchgTile; tile < end; tile += gridDim.x * kXchgTile) {
        const uint32_t nPix = min((uint32_t)kXchgTile, end - tile);
        const uint32_t nWords = (nPix * 3u + 1u) / 2u;                 // 128-bit words; an odd pixel count ends in half a word
        const size_t word0 = (size_t)tile * 3 / 2;
        for (uint32_t k = threadIdx.x; k < nWords; k += blockDim.x) {
            ulonglong2 v[kMaxPeers];
#pragma unroll
            for (int p = 0; p < kMaxPeers; ++p)
                if (p < world) {
                    const ulonglong2* src = reinterpret_cast<const ulonglong2*>(peers.acc[p]) + word0 + k;
                    if (k == nWords - 1 && ((nPix * 3u) & 1u)) v[p] = make_ulonglong2(__ldcv(reinterpret_cast<const unsigned long long*>(src)), 0ull);   // the frame's very last int64
                    else v[p] = __ldcv(src);
                }
            ulonglong2 sum = make_ulonglong2(0ull, 0ull);
#pragma unroll
            for (int p = 0; p < kMaxPeers; ++p)
                if (p < world) {
                    sum.x += v[p].x;
                    sum.y += v[p].y;
                }
            sSum[k] = sum;
        }
        __syncthreads();
        for (uint32_t q = threadIdx.x; q < nPix; q += blockDim.x) {
            const uint32_t i = tile + q;
            const int x = (int)(i % (uint32_t)W), y = (int)(i / (uint32_t)W);
            uchar4 px;
            px.z = to_uchar(accum_to_radiance(sWords[3 * q + 0], invSpp));
            px.y = to_uchar(accum_to_radiance(sWords[3 * q + 1], invSpp));
            px.x = to_uchar(accum_to_radiance(sWords[3 * q + 2], invSpp));
            px.w = (unsigned char)alpha;
            const size_t at = (size_t)(H - y - 1) * W + x;
            if (root >= 0) {
                peers.fb[root][at] = px;
            } else {
                for (int p = 0; p < world; ++p) peers.fb[p][at] = px;
            }
        }
        __syncthreads();
    }
    // completion: every thread's framebuffer stores are made visible system-wide, then the block is counted; the last one signals
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        XchgFlags* mine = peers.flags[rank];
        sLast = atomicAdd(&mine->blocks_done, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (sLast) {
        __threadfence_system();
        if (threadIdx.x == 0) peers.flags[rank]->blocks_done = 0u;   // for the next frame (ordered by the stream)
        if ((int)threadIdx.x < world) st_release_sys(&peers.flags[threadIdx.x]->done[rank], seq);
    }
}

}   // namespace tpt

struct tpt_exchange {
    tpt_context* ctx = nullptr;
    tpt_accum* acc = nullptr;
    int rank = 0, world = 1;
    unsigned char* block = nullptr;          // this rank's shared block: XchgFlags | framebuffer
    tpt::XchgPeers peers{};
    void* opened[2 * tpt::kMaxPeers] = {};   // IPC mappings to close
    int nOpened = 0;
    bool connected = false;
    unsigned int seq = 0;
    unsigned int alpha = 0;
    unsigned long long timeoutNs = 20ull * 1000000000ull;
};

namespace tpt {

// what a rank publishes: TPT_EXCHANGE_HANDLE_BYTES
struct XchgHandle {
    cudaIpcMemHandle_t acc, block;
    unsigned long long accOffset, blockOffset;   // of the pointer inside the allocation the IPC handle names
    int width, height, rank, world;
    int device, pid;
};
static_assert(sizeof(XchgHandle) == TPT_EXCHANGE_HANDLE_BYTES, "tpt.h: TPT_EXCHANGE_HANDLE_BYTES");

// offset of p inside its allocation (an IPC handle names the whole allocation): cuMemGetAddressRange through the runtime's
// driver entry point lookup — the library does not link libcuda
static int allocation_offset(const void* p, unsigned long long* off) {
    using Fn = CUresult (*)(CUdeviceptr*, size_t*, CUdeviceptr);
    static Fn fn = nullptr;
    if (!fn) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult st;
        TPT_CUDA(cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &st));
        TPT_REQUIRE(f != nullptr && st == cudaDriverEntryPointSuccess, "tpt_exchange: cuMemGetAddressRange is not available from this driver");
        fn = reinterpret_cast<Fn>(f);
    }
    CUdeviceptr base = 0;
    size_t size = 0;
    const CUresult r = fn(&base, &size, (CUdeviceptr)(uintptr_t)p);
    TPT_REQUIRE(r == CUDA_SUCCESS, "tpt_exchange: cuMemGetAddressRange failed");
    *off = (unsigned long long)((uintptr_t)p - (uintptr_t)base);
    return TPT_OK;
}

}   // namespace tpt

using namespace tpt;

extern "C" {

int tpt_comm_init_all(int ngpu, tpt_context** out) {
    TPT_REQUIRE(out != nullptr && ngpu >= 1 && ngpu <= kMaxPeers, "tpt_comm_init_all: need 1..16 GPUs and an output array");
    for (int d = 0; d < ngpu; ++d) out[d] = nullptr;
    int count = 0;
    TPT_CUDA(cudaGetDeviceCount(&count));
    if (ngpu > count) {
        set_error("tpt_comm_init_all: " + std::to_string(ngpu) + " GPUs asked for, " + std::to_string(count) + " visible");
        return TPT_ERR_CUDA;
    }
    int rc = TPT_OK;
    for (int d = 0; d < ngpu && rc == TPT_OK; ++d) rc = tpt_context_create(d, &out[d]);
    for (int a = 0; a < ngpu && rc == TPT_OK; ++a) {
        cudaSetDevice(a);
        for (int b = 0; b < ngpu && rc == TPT_OK; ++b) {
            if (a == b) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, a, b);
            if (!can) {
                set_error("tpt_comm_init_all: GPU " + std::to_string(a) + " cannot access GPU " + std::to_string(b) + " as a peer");
                rc = TPT_ERR_CUDA;
                break;
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) rc = cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        }
    }
    if (rc != TPT_OK) {
        const std::string keep = tpt_last_error();
        for (int d = 0; d < ngpu; ++d) {
            if (out[d]) tpt_context_destroy(out[d]);
            out[d] = nullptr;
        }
        set_error(keep);
    }
    return rc;
}

int tpt_accum_allreduce(tpt_accum** acc, int n) {
    TPT_REQUIRE(acc != nullptr && n >= 1 && n <= kMaxPeers, "tpt_accum_allreduce: need 1..16 accumulators");
    for (int r = 0; r < n; ++r) {
        TPT_REQUIRE(acc[r] != nullptr, "tpt_accum_allreduce: NULL accumulator");
        TPT_REQUIRE(acc[r]->width == acc[0]->width && acc[r]->height == acc[0]->height, "tpt_accum_allreduce: accumulators differ in size");
        for (int q = 0; q < r; ++q) TPT_REQUIRE(acc[q] != acc[r], "tpt_accum_allreduce: the same accumulator twice");
    }
    if (n == 1) return TPT_OK;
    // contexts made one by one (tpt_context_create) instead of by tpt_comm_init_all: enable peer access on first use
    static bool peerOk[64][64];
    for (int a = 0; a < n; ++a)
        for (int b = 0; b < n; ++b) {
            const int da = acc[a]->ctx->device, db = acc[b]->ctx->device;
            if (da == db || da >= 64 || db >= 64 || peerOk[da][db]) continue;
            int can = 0;
            TPT_CUDA(cudaDeviceCanAccessPeer(&can, da, db));
            if (!can) {
                set_error("tpt_accum_allreduce: GPU " + std::to_string(da) + " cannot access GPU " + std::to_string(db) + " as a peer");
                return TPT_ERR_CUDA;
            }
            TPT_CUDA(cudaSetDevice(da));
            const cudaError_t e = cudaDeviceEnablePeerAccess(db, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
            peerOk[da][db] = true;
        }
    const size_t total = (size_t)acc[0]->width * acc[0]->height * 3;
    const size_t per = ((total + n - 1) / n + 1) & ~(size_t)1;   // even slice length: slices start 16-byte aligned
    PeerPtrs src;
    for (int k = 0; k < kMaxPeers; ++k) src.p[k] = k < n ? acc[k]->data : nullptr;
    std::vector<cudaEvent_t> ready(n), reduced(n), gathered(n);
    int rc = TPT_OK;
    auto fail = [&](cudaError_t e, const char* what) {
        if (e != cudaSuccess && rc == TPT_OK) rc = cuda_fail(e, what, __FILE__, __LINE__);
        return e != cudaSuccess;
    };
    for (int r = 0; r < n; ++r) {
        cudaSetDevice(acc[r]->ctx->device);
        if (fail(cudaEventCreateWithFlags(&ready[r], cudaEventDisableTiming), "event")) ready[r] = nullptr;
        if (fail(cudaEventCreateWithFlags(&reduced[r], cudaEventDisableTiming), "event")) reduced[r] = nullptr;
        if (fail(cudaEventCreateWithFlags(&gathered[r], cudaEventDisableTiming), "event")) gathered[r] = nullptr;
        if (rc == TPT_OK) fail(cudaEventRecord(ready[r], acc[r]->ctx->stream), "record");   // this GPU's samples are in its buffer
    }
    // 1. reduce-scatter: GPU r sums slice r of every buffer into its own
    for (int r = 0; r < n && rc == TPT_OK; ++r) {
        tpt_context* ctx = acc[r]->ctx;
        cudaSetDevice(ctx->device);
        for (int p = 0; p < n; ++p)
            if (p != r) fail(cudaStreamWaitEvent(ctx->stream, ready[p], 0), "wait");
        const size_t begin = std::min(total, (size_t)r * per), end = std::min(total, begin + per);
        if (end > begin && rc == TPT_OK) {
            const unsigned grid = (unsigned)std::min<size_t>((end - begin + 511) / 512, (size_t)kNumSMs * 8);
            k_reduce_slice<<<grid, 256, 0, ctx->stream>>>(src, n, acc[r]->data, begin, end);
            fail(cudaGetLastError(), "k_reduce_slice");
            ctx->kernel_launches += 1;
        }
        fail(cudaEventRecord(reduced[r], ctx->stream), "record");
    }
    // 2. all-gather: GPU r fetches the reduced slices of the others
    for (int r = 0; r < n && rc == TPT_OK; ++r) {
        tpt_context* ctx = acc[r]->ctx;
        cudaSetDevice(ctx->device);
        for (int p = 0; p < n && rc == TPT_OK; ++p) {
            if (p == r) continue;
            const size_t begin = std::min(total, (size_t)p * per), end = std::min(total, begin + per);
            if (end == begin) continue;
            fail(cudaStreamWaitEvent(ctx->stream, reduced[p], 0), "wait");
            fail(cudaMemcpyPeerAsync(acc[r]->data + begin, ctx->device, acc[p]->data + begin, acc[p]->ctx->device,
                                     (end - begin) * sizeof(unsigned long long), ctx->stream),
                 "cudaMemcpyPeerAsync");
        }
        fail(cudaEventRecord(gathered[r], ctx->stream), "record");
    }
    // 3. nobody may touch its buffer again (next frame's clear) before the peers have read it: every stream waits for all
    for (int r = 0; r < n && rc == TPT_OK; ++r) {
        cudaSetDevice(acc[r]->ctx->device);
        for (int p = 0; p < n; ++p)
            if (p != r) fail(cudaStreamWaitEvent(acc[r]->ctx->stream, gathered[p], 0), "wait");
    }
    if (rc != TPT_OK)   // something was not enqueued: leave no half-ordered work behind
        for (int r = 0; r < n; ++r) {
            cudaSetDevice(acc[r]->ctx->device);
            cudaStreamSynchronize(acc[r]->ctx->stream);
        }
    for (int r = 0; r < n; ++r) {
        cudaSetDevice(acc[r]->ctx->device);
        if (ready[r]) cudaEventDestroy(ready[r]);      // destruction is deferred by the runtime until the event has completed
        if (reduced[r]) cudaEventDestroy(reduced[r]);
        if (gathered[r]) cudaEventDestroy(gathered[r]);
    }
    return rc;
}

// ---- frame exchange: one process per GPU ---------------------------------------------------------------------------------
int tpt_exchange_create(tpt_context* ctx, tpt_accum* acc, int rank, int world, unsigned char alpha, tpt_exchange** out, void* handle_out) {
    TPT_REQUIRE(ctx && acc && out && handle_out, "tpt_exchange_create: NULL argument");
    *out = nullptr;
    TPT_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "tpt_exchange_create: need 0 <= rank < world <= 16");
    TPT_REQUIRE(acc->ctx == ctx, "tpt_exchange_create: the accumulator belongs to another context");
    TPT_REQUIRE(acc->owned, "tpt_exchange_create: the accumulator must be one made by tpt_accum_create (peers map the allocation itself)");
    TPT_REQUIRE((size_t)acc->width * acc->height < ((size_t)1 << 31), "tpt_exchange_create: frame too large");
    TPT_CUDA(cudaSetDevice(ctx->device));
    tpt_exchange* x = new tpt_exchange;
    x->ctx = ctx;
    x->acc = acc;
    x->rank = rank;
    x->world = world;
    x->alpha = alpha;
    if (const char* t = getenv("TPT_EXCHANGE_TIMEOUT_MS")) x->timeoutNs = (unsigned long long)std::max(1ll, atoll(t)) * 1000000ull;
    const size_t bytes = kXchgFbOffset + (size_t)acc->width * acc->height * 4;
    cudaError_t e = cudaMalloc(&x->block, bytes);
    if (e == cudaSuccess) e = cudaMemset(x->block, 0, bytes);
    if (e != cudaSuccess) {
        if (x->block) cudaFree(x->block);
        delete x;
        return cuda_fail(e, "tpt_exchange_create: shared block", __FILE__, __LINE__);
    }
    x->peers.flags[rank] = reinterpret_cast<XchgFlags*>(x->block);
    x->peers.acc[rank] = acc->data;
    x->peers.fb[rank] = reinterpret_cast<uchar4*>(x->block + kXchgFbOffset);
    XchgHandle h;
    memset(&h, 0, sizeof h);
    h.width = acc->width; h.height = acc->height; h.rank = rank; h.world = world;
    h.device = ctx->device; h.pid = (int)getpid();
    int rc = TPT_OK;
    if (world > 1) {
        auto step = [&](cudaError_t err, const char* what) {
            if (err != cudaSuccess && rc == TPT_OK) rc = cuda_fail(err, what, __FILE__, __LINE__);
        };
        step(cudaIpcGetMemHandle(&h.acc, acc->data), "cudaIpcGetMemHandle(accumulator)");
        step(cudaIpcGetMemHandle(&h.block, x->block), "cudaIpcGetMemHandle(shared block)");
        if (rc == TPT_OK) rc = allocation_offset(acc->data, &h.accOffset);
        if (rc == TPT_OK) rc = allocation_offset(x->block, &h.blockOffset);
    } else {
        x->connected = true;
    }
    if (rc != TPT_OK) {
        cudaFree(x->block);
        delete x;
        return rc;
    }
    memcpy(handle_out, &h, sizeof h);
    *out = x;
    return TPT_OK;
}

int tpt_exchange_connect(tpt_exchange* x, const void* all_handles) {
    TPT_REQUIRE(x && all_handles, "tpt_exchange_connect: NULL argument");
    TPT_REQUIRE(!x->connected || x->world == 1, "tpt_exchange_connect: already connected");
    if (x->world == 1) return TPT_OK;
    TPT_CUDA(cudaSetDevice(x->ctx->device));
    const unsigned char* raw = static_cast<const unsigned char*>(all_handles);
    for (int p = 0; p < x->world; ++p) {
        XchgHandle h;
        memcpy(&h, raw + (size_t)p * sizeof h, sizeof h);
        TPT_REQUIRE(h.rank == p && h.world == x->world, "tpt_exchange_connect: handle p must come from rank p of the same world");
        TPT_REQUIRE(h.width == x->acc->width && h.height == x->acc->height, "tpt_exchange_connect: the ranks' frames differ in size");
        if (p == x->rank) continue;
        TPT_REQUIRE(h.pid != (int)getpid(), "tpt_exchange_connect: two ranks in one process — use tpt_accum_allreduce there");
        void *a = nullptr, *b = nullptr;
        TPT_CUDA(cudaIpcOpenMemHandle(&a, h.acc, cudaIpcMemLazyEnablePeerAccess));
        x->opened[x->nOpened++] = a;
        TPT_CUDA(cudaIpcOpenMemHandle(&b, h.block, cudaIpcMemLazyEnablePeerAccess));
        x->opened[x->nOpened++] = b;
        unsigned char* blk = static_cast<unsigned char*>(b) + h.blockOffset;
        x->peers.acc[p] = reinterpret_cast<const unsigned long long*>(static_cast<unsigned char*>(a) + h.accOffset);
        x->peers.flags[p] = reinterpret_cast<XchgFlags*>(blk);
        x->peers.fb[p] = reinterpret_cast<uchar4*>(blk + kXchgFbOffset);
    }
    x->connected = true;
    return TPT_OK;
}

int tpt_exchange_resolve(tpt_exchange* x, int spp_total, int root) {
    TPT_REQUIRE(x != nullptr, "tpt_exchange_resolve: NULL exchange");
    TPT_REQUIRE(x->connected, "tpt_exchange_resolve: tpt_exchange_connect has not run");
    TPT_REQUIRE(spp_total >= 1 && root >= -1 && root < x->world, "tpt_exchange_resolve: need spp_total >= 1 and -1 <= root < world");
    tpt_context* ctx = x->ctx;
    TPT_CUDA(cudaSetDevice(ctx->device));
    const unsigned int seq = ++x->seq;
    const uint32_t total = (uint32_t)((size_t)x->acc->width * x->acc->height);
    const uint32_t per = ((total + x->world - 1) / x->world + 1u) & ~1u;   // even: every rank's part starts on a 16-byte boundary
    const uint32_t begin = std::min(total, (uint32_t)x->rank * per), end = std::min(total, begin + per);
    if (x->world > 1) k_xchg_ready<<<1, kMaxPeers, 0, ctx->stream>>>(x->peers, x->rank, x->world, seq, x->timeoutNs);
    const unsigned grid = std::max(1u, std::min((end - begin + kXchgTile - 1) / kXchgTile, (unsigned)kNumSMs * 8));
    k_xchg_resolve<<<grid, 256, 0, ctx->stream>>>(x->peers, x->rank, x->world, root, seq, begin, end, x->acc->width, x->acc->height,
                                                  1.0f / (float)spp_total, x->alpha);
    if (x->world > 1) k_xchg_wait_done<<<1, kMaxPeers, 0, ctx->stream>>>(x->peers, x->rank, x->world, seq, x->timeoutNs);
    TPT_CUDA(cudaGetLastError());
    ctx->kernel_launches += x->world > 1 ? 3 : 1;
    return TPT_OK;
}

void* tpt_exchange_framebuffer(tpt_exchange* x) { return x ? static_cast<void*>(x->block + kXchgFbOffset) : nullptr; }

int tpt_exchange_read_framebuffer(tpt_exchange* x, uint8_t* host_fb) {
    TPT_REQUIRE(x && host_fb, "tpt_exchange_read_framebuffer: NULL argument");
    TPT_CUDA(cudaSetDevice(x->ctx->device));
    TPT_CUDA(cudaStreamSynchronize(x->ctx->stream));
    TPT_CUDA(cudaMemcpy(host_fb, x->block + kXchgFbOffset, (size_t)x->acc->width * x->acc->height * 4, cudaMemcpyDeviceToHost));
    return TPT_OK;
}

int tpt_exchange_status(tpt_exchange* x) {
    TPT_REQUIRE(x != nullptr, "tpt_exchange_status: NULL exchange");
    TPT_CUDA(cudaSetDevice(x->ctx->device));
    TPT_CUDA(cudaStreamSynchronize(x->ctx->stream));
    XchgFlags f;
    TPT_CUDA(cudaMemcpy(&f, x->block, sizeof f, cudaMemcpyDeviceToHost));
    if (f.error) {
        set_error("tpt_exchange: a wait for another rank timed out (rank " + std::to_string(x->rank) + " of " + std::to_string(x->world) +
                  ", frame " + std::to_string(x->seq) + "); the framebuffer is not valid");
        return TPT_ERR_CUDA;
    }
    return TPT_OK;
}

int tpt_exchange_destroy(tpt_exchange* x) {
    if (!x) return TPT_OK;
    cudaSetDevice(x->ctx->device);
    cudaStreamSynchronize(x->ctx->stream);
    for (int i = 0; i < x->nOpened; ++i) cudaIpcCloseMemHandle(x->opened[i]);
    cudaFree(x->block);
    delete x;
    return TPT_OK;
}

}   // extern "C"
