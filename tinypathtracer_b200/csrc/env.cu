// env.cu — builds the environment-map alias table (see env.cuh) and the batch sampling entry tpt_env_sample.
//
// Replaces computeEnvLightPDF + thrust::inclusive_scan + initEnvLight of the reference (src/env_light.cu:10-36,50-53).
// Texel weights are computed on the device; Vose's O(n) table construction runs once per environment map on the host
// (it is a sequential work-list algorithm and not on any per-frame path), the tables live on the device.
#include <vector>

#include "env.cuh"
#include "scene.cuh"

namespace tpt {

// weight(texel) = luminance(rgb) * sin(theta of the texel's centre row); row 0 = bottom <-> v = 0 <-> theta = pi
__global__ void __launch_bounds__(256) k_env_weights(const uchar4* __restrict__ env, const float4* __restrict__ envf, int w, int h,
                                                     float* __restrict__ weight) {
    const size_t n = (size_t)w * h;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / (size_t)w);
        float3 c;
        if (envf != nullptr) {
            const float4 t = envf[i];
            c = f3(t.x, t.y, t.z);
        } else {
            const uchar4 t = env[i];
            c = f3((float)t.x, (float)t.y, (float)t.z);
        }
        const float lum = 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z;
        weight[i] = lum * sinf((1.0f - ((float)y + 0.5f) / (float)h) * kPi);
    }
}

__global__ void __launch_bounds__(256) k_env_sample(EnvSampler es, size_t n, uint64_t seedKey, float4* __restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint64_t key = rng_path_key(seedKey, (uint32_t)(i >> 32), (uint32_t)i);
        float u1, u2, u3, u4;
        rng_pair(key, 1u, u1, u2);
        rng_pair(key, 2u, u3, u4);
        float pdf;
        const float3 d = env_sample_dir(es, (uint32_t)(rng_bits(key, 1u) >> 32), u4, u2, u3, pdf);
        out[i] = make_float4(d.x, d.y, d.z, pdf);
    }
}

}   // namespace tpt

using namespace tpt;

int tpt_env_free_sampler(tpt_context* ctx) {
    cudaFree(ctx->env_prob);
    cudaFree(ctx->env_alias);
    cudaFree(ctx->env_mass);
    ctx->env_prob = ctx->env_mass = nullptr;
    ctx->env_alias = nullptr;
    return TPT_OK;
}

// Vose's alias method: buckets of height 1/n, each holding at most two texels.
int tpt_env_build_sampler(tpt_context* ctx) {
    if (ctx->env_prob) return TPT_OK;
    TPT_REQUIRE(ctx->env != nullptr || ctx->envf != nullptr, "environment importance sampling needs an environment map (tpt_env_set_rgba8 / tpt_env_set_rgb32f)");
    const size_t n = (size_t)ctx->env_w * ctx->env_h;
    float* d_weight = nullptr;
    TPT_CUDA(cudaMalloc(&d_weight, n * sizeof(float)));
    k_env_weights<<<kNumSMs * 8, 256, 0, ctx->stream>>>(ctx->env, ctx->envf, ctx->env_w, ctx->env_h, d_weight);
    ctx->kernel_launches += 1;
    std::vector<float> wgt(n);
    cudaError_t e = cudaMemcpyAsync(wgt.data(), d_weight, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_weight);
    TPT_CUDA(e);
    double sum = 0.0;
    for (size_t i = 0; i < n; ++i) sum += wgt[i];
    std::vector<float> mass(n), prob(n);
    std::vector<uint32_t> alias(n);
    if (!(sum > 0.0)) {   // black map: uniform over the sphere's texels by solid angle is pointless, fall back to uniform texels
        for (size_t i = 0; i < n; ++i) mass[i] = 1.0f / (float)n;
    } else {
        for (size_t i = 0; i < n; ++i) mass[i] = (float)(wgt[i] / sum);
    }
    std::vector<double> scaled(n);
    std::vector<uint32_t> small, large;
    small.reserve(n);
    large.reserve(n);
    for (size_t i = 0; i < n; ++i) {
        scaled[i] = (double)mass[i] * (double)n;
        (scaled[i] < 1.0 ? small : large).push_back((uint32_t)i);
    }
    while (!small.empty() && !large.empty()) {
        const uint32_t s = small.back(), l = large.back();
        small.pop_back();
        prob[s] = (float)scaled[s];
        alias[s] = l;
        scaled[l] = (scaled[l] + scaled[s]) - 1.0;
        if (scaled[l] < 1.0) {
            large.pop_back();
            small.push_back(l);
        }
    }
    for (uint32_t i : large) { prob[i] = 1.0f; alias[i] = i; }
    for (uint32_t i : small) { prob[i] = 1.0f; alias[i] = i; }   // rounding leftovers
    TPT_CUDA(cudaMalloc(&ctx->env_prob, n * sizeof(float)));
    TPT_CUDA(cudaMalloc(&ctx->env_alias, n * sizeof(uint32_t)));
    TPT_CUDA(cudaMalloc(&ctx->env_mass, n * sizeof(float)));
    TPT_CUDA(cudaMemcpyAsync(ctx->env_prob, prob.data(), n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    TPT_CUDA(cudaMemcpyAsync(ctx->env_alias, alias.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    TPT_CUDA(cudaMemcpyAsync(ctx->env_mass, mass.data(), n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    return TPT_OK;
}

int tpt_env_sample_impl(tpt_context* ctx, size_t n, uint64_t seed, float* host_dirs_pdf) {
    TPT_CHECK(tpt_env_build_sampler(ctx));
    float4* d = nullptr;
    TPT_CUDA(cudaMalloc(&d, n * sizeof(float4)));
    const EnvSampler es{ctx->env_prob, ctx->env_alias, ctx->env_mass, ctx->env_w, ctx->env_h};
    k_env_sample<<<kNumSMs * 8, 256, 0, ctx->stream>>>(es, n, rng_seed_key(seed), d);
    ctx->kernel_launches += 1;
    cudaError_t e = cudaMemcpyAsync(host_dirs_pdf, d, n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    TPT_CUDA(e);
    return TPT_OK;
}
