// comm.cu — the one exchange step of the multi-GPU path, for a host that drives all GPUs from ONE process (as the
// reference drives its single GPU: one host thread, src/path_tracer.cu:491-554): the sum of the per-GPU accumulation
// buffers.  No library collective: every GPU reduces its own 1/n slice by reading the peers' buffers
// directly over NVLink (peer loads in k_reduce_slice), then collects the other reduced slices with peer copies.
// The buffers are int64 fixed point, so the sum is exact and the result does not depend on n or on the order.
// (One process PER GPU — bench.py, torch.distributed — wraps the accumulator in a tensor and uses NCCL's all-reduce
// instead: tpt_accum_wrap / tpt_accum_device_ptr.)
#include <algorithm>
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

}   // extern "C"
