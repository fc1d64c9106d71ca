// capi.cu — the extern "C" boundary declared in include/tpt.h.
//
// Host-side orchestration only: argument checks, device memory ownership, host<->device staging and
// kernel sequencing.  There is no CPU implementation of any entry point: when no CUDA device is
// usable, tpt_context_create fails and nothing else can be called.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "scene.cuh"

namespace tpt {

static thread_local std::string g_last_error;

void set_error(const std::string& msg) { g_last_error = msg; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    g_last_error = std::string("CUDA error ") + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ") at " + file + ":" +
                   std::to_string(line) + " in " + what;
    return TPT_ERR_CUDA;
}

}   // namespace tpt

using namespace tpt;

// implemented in bvh.cu / render.cu
int tpt_scene_build_tables(tpt_scene* s);
int tpt_bvh_build_impl(tpt_scene* s);
int tpt_bvh_refit_impl(tpt_scene* s);
int tpt_bvh_stats_impl(tpt_scene* s, tpt_bvh_stats* stats);
int tpt_scene_ensure_normals(tpt_scene* s);   // bvh.cu
int tpt_bvh_export_reference_impl(tpt_scene* s, void* out, size_t nNodes);
int tpt_primary_rays_impl(tpt_context* ctx, const tpt_camera* cam, int W, int H, int jitter, uint64_t seed, tpt_ray* d_rays);
int tpt_trace_rays_impl(tpt_scene* s, const tpt_ray* d_rays, size_t n, tpt_hit* d_hits, int anyHit, int counted);
int tpt_render_impl(tpt_scene* s, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc);
int tpt_accum_add_impl(tpt_accum* dst, const tpt_accum* src);
int tpt_accum_to_f32_impl(tpt_accum* acc, float* d_out);
int tpt_resolve_impl(tpt_accum* acc, int sppTotal, uint8_t* d_fb);
int tpt_env_build_sampler(tpt_context* ctx);
int tpt_env_free_sampler(tpt_context* ctx);
int tpt_env_sample_impl(tpt_context* ctx, size_t n, uint64_t seed, float* host_dirs_pdf);

namespace {

template <typename T>
int dev_alloc(T** p, size_t count) {
    *p = nullptr;
    if (count == 0) count = 1;
    TPT_CUDA(cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T)));
    return TPT_OK;
}
template <typename T>
int upload(tpt_context* ctx, T* dst, const T* src, size_t count) {
    if (count == 0) return TPT_OK;
    TPT_CUDA(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    return TPT_OK;
}
int bind(tpt_context* ctx) {
    TPT_CUDA(cudaSetDevice(ctx->device));
    return TPT_OK;
}
int ensure_io(tpt_context* ctx, size_t n) {
    if (n <= ctx->io_capacity) return TPT_OK;
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->io_rays) cudaFree(ctx->io_rays);
    if (ctx->io_hits) cudaFree(ctx->io_hits);
    ctx->io_rays = ctx->io_hits = nullptr;
    ctx->io_capacity = 0;
    TPT_CUDA(cudaMalloc(&ctx->io_rays, n * sizeof(tpt_ray)));
    TPT_CUDA(cudaMalloc(&ctx->io_hits, n * sizeof(tpt_hit)));
    ctx->io_capacity = n;
    return TPT_OK;
}

}   // namespace

extern "C" {

const char* tpt_last_error(void) { return g_last_error.c_str(); }
int tpt_abi_version(void) { return TPT_ABI_VERSION; }

int tpt_context_create(int device, tpt_context** out) {
    TPT_REQUIRE(out != nullptr, "tpt_context_create: out is NULL");
    *out = nullptr;
    int count = 0;
    TPT_CUDA(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) {
        set_error("tpt_context_create: no such CUDA device " + std::to_string(device) + " (" + std::to_string(count) + " visible); there is no CPU fallback");
        return TPT_ERR_CUDA;
    }
    TPT_CUDA(cudaSetDevice(device));
    tpt_context* ctx = new (std::nothrow) tpt_context();
    if (!ctx) return TPT_ERR_NOMEM;
    ctx->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&ctx->d_counters, sizeof(DeviceCounters));
    if (e == cudaSuccess) e = cudaMemset(ctx->d_counters, 0, sizeof(DeviceCounters));
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
    if (e != cudaSuccess) {
        delete ctx;
        return cuda_fail(e, "context setup", __FILE__, __LINE__);
    }
    ctx->own_stream = true;
    *out = ctx;
    return TPT_OK;
}

int tpt_context_destroy(tpt_context* ctx) {
    if (!ctx) return TPT_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->env);
    cudaFree(ctx->envf);
    tpt_env_free_sampler(ctx);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->q_ext[0]);
    cudaFree(ctx->q_ext[1]);
    cudaFree(ctx->q_sec[0]);
    cudaFree(ctx->q_sec[1]);
    cudaFree(ctx->levels);
    cudaFree(ctx->trace_cursor);
    cudaFree(ctx->d_tile_live);
    if (ctx->io_in) cudaStreamDestroy(ctx->io_in);
    if (ctx->io_out) cudaStreamDestroy(ctx->io_out);
    for (cudaEvent_t e : ctx->io_events) cudaEventDestroy(e);
    cudaFree(ctx->io_rays);
    cudaFree(ctx->io_hits);
    for (cudaEvent_t e : ctx->stage_events) cudaEventDestroy(e);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return TPT_OK;
}

int tpt_context_synchronize(tpt_context* ctx) {
    TPT_REQUIRE(ctx, "tpt_context_synchronize: ctx is NULL");
    TPT_CHECK(bind(ctx));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    return TPT_OK;
}

int tpt_context_set_stream(tpt_context* ctx, void* cuda_stream) {
    TPT_REQUIRE(ctx, "tpt_context_set_stream: ctx is NULL");
    TPT_CHECK(bind(ctx));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
    ctx->own_stream = false;
    return TPT_OK;
}

int tpt_get_counters(tpt_context* ctx, tpt_counters* out) {
    TPT_REQUIRE(ctx && out, "tpt_get_counters: NULL argument");
    TPT_CHECK(bind(ctx));
    DeviceCounters h;
    TPT_CUDA(cudaMemcpyAsync(&h, ctx->d_counters, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    out->rays_primary = h.rays_primary;
    out->rays_extension = h.rays_extension;
    out->rays_direct = h.rays_direct;
    out->rays_shadow = h.rays_shadow;
    out->node_visits = h.node_visits;
    out->tri_tests = h.tri_tests;
    out->samples = ctx->samples;
    out->dropped_nonfinite = h.dropped_nonfinite;
    out->kernel_launches = ctx->kernel_launches;
    out->rays_culled = h.rays_culled;
    out->rays_candidates = h.rays_candidates;
    out->queue_overflow = h.queue_overflow;
    out->rays_queued = h.rays_queued;
    return TPT_OK;
}

int tpt_reset_counters(tpt_context* ctx) {
    TPT_REQUIRE(ctx, "tpt_reset_counters: ctx is NULL");
    TPT_CHECK(bind(ctx));
    TPT_CUDA(cudaMemsetAsync(ctx->d_counters, 0, sizeof(DeviceCounters), ctx->stream));
    ctx->samples = 0;
    ctx->kernel_launches = 0;
    for (int k = 0; k < 4; ++k) {
        ctx->stage_ms[k] = 0.0;
        ctx->stage_launches[k] = 0;
    }
    return TPT_OK;
}

int tpt_get_stage_times(tpt_context* ctx, tpt_stage_times* out) {
    TPT_REQUIRE(ctx && out, "tpt_get_stage_times: NULL argument");
    for (int k = 0; k < 4; ++k) {
        out->ms[k] = ctx->stage_ms[k];
        out->launches[k] = ctx->stage_launches[k];
    }
    return TPT_OK;
}

int tpt_scene_destroy(tpt_scene* s) {
    if (!s) return TPT_OK;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    cudaFree(s->indices); cudaFree(s->obj_verts); cudaFree(s->obj_norms); cudaFree(s->vert_obj); cudaFree(s->face_mtl);
    cudaFree(s->objects); cudaFree(s->vert_trans); cudaFree(s->normal_trans); cudaFree(s->mtl_table); cudaFree(s->lights);
    cudaFree(s->world_verts); cudaFree(s->world_norms); cudaFree(s->keys[0]); cudaFree(s->keys[1]); cudaFree(s->vals[0]);
    cudaFree(s->vals[1]); cudaFree(s->sort_tmp); cudaFree(s->parent_int); cudaFree(s->parent_leaf); cudaFree(s->flags); cudaFree(s->merge_slots);
    cudaFree(s->bounds); cudaFree(s->nodes); cudaFree(s->nodes4); cudaFree(s->tri_geom); cudaFree(s->tri_norm);
    for (int k = 0; k < 4; ++k) {
        if (s->h_mtl[k]) cudaFreeHost(s->h_mtl[k]);
        if (s->h_mtl_free[k]) cudaEventDestroy(s->h_mtl_free[k]);
    }
    delete s;
    return TPT_OK;
}

// argument checks shared by tpt_scene_create and tpt_scene_update
static int check_scene_desc(const tpt_scene_desc* d, const char* who) {
    const std::string w(who);
    TPT_REQUIRE(d->n_indices >= 3 && d->n_indices % 3 == 0, w + ": n_indices must be a positive multiple of 3");
    TPT_REQUIRE(d->n_indices / 3 < (size_t)1 << 30, w + ": too many faces");
    TPT_REQUIRE(d->n_vertices > 0 && d->indices && d->vertices && d->normals, w + ": geometry arrays missing");
    TPT_REQUIRE(d->n_objects > 0 && d->objects && d->vert_trans && d->normal_trans, w + ": object table missing");
    TPT_REQUIRE(d->n_materials > 0 && d->materials, w + ": at least one material is required");
    TPT_REQUIRE(d->n_lights == 0 || d->lights, w + ": lights pointer is NULL");
    const size_t F = d->n_indices / 3;
    for (size_t i = 0; i < d->n_indices; ++i) TPT_REQUIRE(d->indices[i] < d->n_vertices, w + ": vertex index out of range");
    TPT_REQUIRE(d->objects[0].begin == 0, w + ": first object must begin at face 0");
    for (size_t o = 0; o < d->n_objects; ++o) {
        TPT_REQUIRE(d->objects[o].mtl_idx >= 0 && (size_t)d->objects[o].mtl_idx < d->n_materials, w + ": material index out of range");
        TPT_REQUIRE(o == 0 || d->objects[o].begin >= d->objects[o - 1].begin, w + ": object face ranges must ascend");
        TPT_REQUIRE((size_t)d->objects[o].begin <= F, w + ": object begin beyond face count");
    }
    return TPT_OK;
}

// host arrays -> the scene's device arrays (+ the per-scene tables derived from them); no device allocation.  The sources may be
// pageable (cudaMemcpyAsync then returns once the buffer has been staged, and may first wait for the stream) or page-locked (the
// copies are asynchronous: the caller keeps the arrays alive until the stream has passed them — bench.py's e2e does this).
static int upload_scene(tpt_scene* s, const tpt_scene_desc* d) {
    tpt_context* ctx = s->ctx;
    const size_t F = s->n_faces, V = s->n_vertices, O = s->n_objects;
    s->any_emissive = s->any_specular = false;
    for (size_t m = 0; m < d->n_materials; ++m) s->any_emissive = s->any_emissive || d->materials[m].emission > 0.0f;
    for (size_t m = 0; m < d->n_materials; ++m) s->any_specular = s->any_specular || d->materials[m].eta > 0.0f || d->materials[m].metallic > 0.0f;
    const unsigned slot = s->h_mtl_next++ & 3u;
    if (!s->h_mtl[slot]) {
        TPT_CUDA(cudaMallocHost(reinterpret_cast<void**>(&s->h_mtl[slot]), 2 * s->n_materials * sizeof(float4)));
        TPT_CUDA(cudaEventCreateWithFlags(&s->h_mtl_free[slot], cudaEventDisableTiming));
    } else {
        TPT_CUDA(cudaEventSynchronize(s->h_mtl_free[slot]));   // the upload four updates ago has read this buffer
    }
    float4* mtl = s->h_mtl[slot];
    for (size_t m = 0; m < d->n_materials; ++m) {
        const tpt_material& M = d->materials[m];
        mtl[2 * m] = make_float4(M.base_color[0], M.base_color[1], M.base_color[2], M.emission);
        mtl[2 * m + 1] = make_float4(M.eta, M.metallic, M.unused[2] /* roughness (include/material.h:113) */, 0.f);
    }
    TPT_CHECK(upload(ctx, s->indices, d->indices, 3 * F));
    TPT_CHECK(upload(ctx, s->obj_verts, d->vertices, 3 * V));
    TPT_CHECK(upload(ctx, s->obj_norms, d->normals, 3 * V));
    TPT_CHECK(upload(ctx, s->objects, d->objects, O));
    TPT_CHECK(upload(ctx, s->vert_trans, d->vert_trans, 16 * O));
    TPT_CHECK(upload(ctx, s->normal_trans, d->normal_trans, 16 * O));
    TPT_CHECK(upload(ctx, s->mtl_table, mtl, 2 * d->n_materials));
    TPT_CUDA(cudaEventRecord(s->h_mtl_free[slot], ctx->stream));
    TPT_CHECK(upload(ctx, s->lights, d->lights, d->n_lights));
    return tpt_scene_build_tables(s);
}

int tpt_scene_create(tpt_context* ctx, const tpt_scene_desc* d, tpt_scene** out) {
    TPT_REQUIRE(ctx && d && out, "tpt_scene_create: NULL argument");
    *out = nullptr;
    TPT_CHECK(check_scene_desc(d, "tpt_scene_create"));
    const size_t F = d->n_indices / 3;
    TPT_CHECK(bind(ctx));
    tpt_scene* s = new (std::nothrow) tpt_scene();
    if (!s) return TPT_ERR_NOMEM;
    s->ctx = ctx;
    s->n_faces = F;
    s->n_vertices = d->n_vertices;
    s->n_objects = d->n_objects;
    s->n_materials = d->n_materials;
    s->n_lights = d->n_lights;
    const size_t V = d->n_vertices, O = d->n_objects;
    int rc = TPT_OK;
    auto A = [&](int r) { if (rc == TPT_OK) rc = r; };
    A(dev_alloc(&s->indices, 3 * F)); A(dev_alloc(&s->obj_verts, 3 * V)); A(dev_alloc(&s->obj_norms, 3 * V));
    A(dev_alloc(&s->vert_obj, V)); A(dev_alloc(&s->face_mtl, F)); A(dev_alloc(&s->objects, O));
    A(dev_alloc(&s->vert_trans, 16 * O)); A(dev_alloc(&s->normal_trans, 16 * O)); A(dev_alloc(&s->mtl_table, 2 * d->n_materials));
    A(dev_alloc(&s->lights, d->n_lights)); A(dev_alloc(&s->world_verts, 3 * V)); A(dev_alloc(&s->world_norms, 3 * V));
    A(dev_alloc(&s->keys[0], F)); A(dev_alloc(&s->keys[1], F)); A(dev_alloc(&s->vals[0], F)); A(dev_alloc(&s->vals[1], F));
    A(dev_alloc(&s->parent_int, F)); A(dev_alloc(&s->parent_leaf, F)); A(dev_alloc(&s->flags, F)); A(dev_alloc(&s->bounds, 8));
    A(dev_alloc(&s->nodes, 4 * F)); A(dev_alloc(&s->tri_geom, 3 * F)); A(dev_alloc(&s->tri_norm, 3 * F));
    A(dev_alloc(&s->merge_slots, 4 * F));
    // 4-wide compressed nodes beside the binary ones (64 B per internal node; traverse.cuh) — OPT-IN: TPT_WIDE_MIN_FACES=<n> builds and
    // walks them for scenes of n triangles and more (the tests use 2).  Measured on the 10 M-triangle soup they halve the node visits
    // (72.9 -> 36.7 per incoherent ray) but not the time (1 274 vs 1 279 Mrays/s; coherent primary rays 1 893 vs 2 270): a wide visit
    // costs ~3.4x the instructions of a binary one and pushes up to three stack entries, so the walk turns from L1-bound to
    // issue-bound at the same speed, and the collapse adds 0.35 ms to a 10 M build.  Off by default until that is fixed.
    static const size_t wideMinFaces = getenv("TPT_WIDE_MIN_FACES") ? (size_t)atoll(getenv("TPT_WIDE_MIN_FACES")) : ~(size_t)0;
    if (F >= wideMinFaces && F >= 2) A(dev_alloc(&s->nodes4, 4 * F));
    if (rc == TPT_OK) rc = upload_scene(s, d);
    if (rc == TPT_OK) {
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "scene upload", __FILE__, __LINE__);
    }
    if (rc != TPT_OK) {
        std::string keep = g_last_error;
        tpt_scene_destroy(s);
        g_last_error = keep;
        return rc;
    }
    *out = s;
    return TPT_OK;
}

int tpt_scene_update(tpt_scene* s, const tpt_scene_desc* d) {
    TPT_REQUIRE(s && d, "tpt_scene_update: NULL argument");
    TPT_CHECK(check_scene_desc(d, "tpt_scene_update"));
    TPT_REQUIRE(d->n_indices / 3 == s->n_faces && d->n_vertices == s->n_vertices && d->n_objects == s->n_objects &&
                    d->n_materials == s->n_materials && d->n_lights == s->n_lights,
                "tpt_scene_update: array sizes differ from the scene's (faces, vertices, objects, materials and lights must match "
                "tpt_scene_create's; create a new scene otherwise)");
    TPT_CHECK(bind(s->ctx));
    s->built = false;
    s->topology = false;   // new geometry: the hierarchy of the last build says nothing about it
    return upload_scene(s, d);
}

int tpt_scene_set_transforms(tpt_scene* s, const float* vert_trans, const float* normal_trans) {
    TPT_REQUIRE(s && vert_trans && normal_trans, "tpt_scene_set_transforms: NULL argument");
    TPT_CHECK(bind(s->ctx));
    TPT_CHECK(upload(s->ctx, s->vert_trans, vert_trans, 16 * s->n_objects));
    TPT_CHECK(upload(s->ctx, s->normal_trans, normal_trans, 16 * s->n_objects));
    TPT_CUDA(cudaStreamSynchronize(s->ctx->stream));
    s->built = false;
    return TPT_OK;
}

int tpt_bvh_build(tpt_scene* s, tpt_bvh_stats* stats) {
    TPT_REQUIRE(s, "tpt_bvh_build: scene is NULL");
    tpt_context* ctx = s->ctx;
    TPT_CHECK(bind(ctx));
    if (stats) TPT_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
    TPT_CHECK(tpt_bvh_build_impl(s));
    if (stats) {
        TPT_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
        TPT_CUDA(cudaEventSynchronize(ctx->ev1));
        memset(stats, 0, sizeof *stats);
        TPT_CHECK(tpt_bvh_stats_impl(s, stats));
        TPT_CUDA(cudaEventElapsedTime(&stats->build_ms, ctx->ev0, ctx->ev1));
    }
    return TPT_OK;
}

int tpt_bvh_build_from_world(tpt_context* ctx, const float* d_verts, size_t n_vertices, const uint32_t* d_indices, size_t n_faces, tpt_scene** scene,
                             tpt_bvh_stats* stats) {
    TPT_REQUIRE(ctx && d_verts && d_indices && scene, "tpt_bvh_build_from_world: NULL argument");
    TPT_REQUIRE(n_vertices > 0 && n_faces > 0 && n_faces < (size_t)1 << 30, "tpt_bvh_build_from_world: bad sizes");
    TPT_CHECK(bind(ctx));
    if (*scene == nullptr) {
        // a one-object scene of the right size: identity matrices (M * v == v exactly), one default material, zero normals
        std::vector<uint32_t> idx(3 * n_faces, 0u);
        std::vector<float> zeros(3 * n_vertices, 0.0f);
        const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        tpt_material mtl;
        std::memset(&mtl, 0, sizeof mtl);
        mtl.base_color[0] = mtl.base_color[1] = mtl.base_color[2] = 0.8f;
        const tpt_mtl_interval obj = {0, 0};
        tpt_scene_desc d;
        std::memset(&d, 0, sizeof d);
        d.indices = idx.data();
        d.n_indices = idx.size();
        d.vertices = zeros.data();
        d.normals = zeros.data();
        d.n_vertices = n_vertices;
        d.materials = &mtl;
        d.n_materials = 1;
        d.objects = &obj;
        d.n_objects = 1;
        d.vert_trans = ident;
        d.normal_trans = ident;
        TPT_CHECK(tpt_scene_create(ctx, &d, scene));
    }
    tpt_scene* s = *scene;
    TPT_REQUIRE(s->ctx == ctx && s->n_faces == n_faces && s->n_vertices == n_vertices && s->n_objects == 1,
                "tpt_bvh_build_from_world: the scene was not made by this call for these sizes");
    TPT_CUDA(cudaMemcpyAsync(s->indices, d_indices, 3 * n_faces * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    TPT_CUDA(cudaMemcpyAsync(s->obj_verts, d_verts, 3 * n_vertices * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    TPT_CHECK(tpt_scene_build_tables(s));   // which vertices the new indices reference
    s->topology = false;
    return tpt_bvh_build(s, stats);
}

int tpt_bvh_refit(tpt_scene* s, tpt_bvh_stats* stats) {
    TPT_REQUIRE(s, "tpt_bvh_refit: scene is NULL");
    if (!s->topology) {
        set_error("tpt_bvh_refit: no hierarchy to reuse, call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    tpt_context* ctx = s->ctx;
    TPT_CHECK(bind(ctx));
    if (stats) TPT_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
    TPT_CHECK(tpt_bvh_refit_impl(s));
    if (stats) {
        TPT_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
        TPT_CUDA(cudaEventSynchronize(ctx->ev1));
        memset(stats, 0, sizeof *stats);
        TPT_CHECK(tpt_bvh_stats_impl(s, stats));
        TPT_CUDA(cudaEventElapsedTime(&stats->build_ms, ctx->ev0, ctx->ev1));
    }
    return TPT_OK;
}

int tpt_bvh_export_reference_nodes(tpt_scene* s, void* out_nodes36, size_t n_nodes) {
    TPT_REQUIRE(s && out_nodes36, "tpt_bvh_export_reference_nodes: NULL argument");
    if (!s->built) {
        set_error("tpt_bvh_export_reference_nodes: call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    TPT_REQUIRE(n_nodes == 2 * s->n_faces - 1, "tpt_bvh_export_reference_nodes: n_nodes must be 2F-1");
    TPT_CHECK(bind(s->ctx));
    return tpt_bvh_export_reference_impl(s, out_nodes36, n_nodes);
}

int tpt_bvh_export_order(tpt_scene* s, uint64_t* keys, uint32_t* faces) {
    TPT_REQUIRE(s, "tpt_bvh_export_order: scene is NULL");
    if (!s->built || !s->topology) {
        set_error("tpt_bvh_export_order: call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    TPT_CHECK(bind(s->ctx));
    if (keys) TPT_CUDA(cudaMemcpyAsync(keys, s->keys[1], s->n_faces * sizeof(uint64_t), cudaMemcpyDeviceToHost, s->ctx->stream));
    if (faces) TPT_CUDA(cudaMemcpyAsync(faces, s->vals[1], s->n_faces * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(s->ctx->stream));
    return TPT_OK;
}

int tpt_scene_export_world(tpt_scene* s, float* wv, float* wn) {
    TPT_REQUIRE(s, "tpt_scene_export_world: scene is NULL");
    if (!s->built) {
        set_error("tpt_scene_export_world: call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    TPT_CHECK(bind(s->ctx));
    if (wn) TPT_CHECK(tpt_scene_ensure_normals(s));   // large scenes transform their normals on first use
    if (wv) TPT_CUDA(cudaMemcpyAsync(wv, s->world_verts, 3 * s->n_vertices * sizeof(float), cudaMemcpyDeviceToHost, s->ctx->stream));
    if (wn) TPT_CUDA(cudaMemcpyAsync(wn, s->world_norms, 3 * s->n_vertices * sizeof(float), cudaMemcpyDeviceToHost, s->ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(s->ctx->stream));
    return TPT_OK;
}

static int env_drop(tpt_context* ctx) {
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->env) cudaFree(ctx->env);
    if (ctx->envf) cudaFree(ctx->envf);
    ctx->env = nullptr;
    ctx->envf = nullptr;
    ctx->env_w = ctx->env_h = 0;
    tpt_env_free_sampler(ctx);
    return TPT_OK;
}

int tpt_env_set_rgba8(tpt_context* ctx, const uint8_t* rgba8, int w, int h) {
    TPT_REQUIRE(ctx && rgba8, "tpt_env_set_rgba8: NULL argument");
    TPT_REQUIRE(w > 0 && h > 0, "tpt_env_set_rgba8: bad size");
    TPT_CHECK(bind(ctx));
    TPT_CHECK(env_drop(ctx));
    TPT_CUDA(cudaMalloc(&ctx->env, (size_t)w * h * 4));
    TPT_CUDA(cudaMemcpyAsync(ctx->env, rgba8, (size_t)w * h * 4, cudaMemcpyHostToDevice, ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->env_w = w;
    ctx->env_h = h;
    return TPT_OK;
}

int tpt_env_set_rgb32f(tpt_context* ctx, const float* rgb, int w, int h) {
    TPT_REQUIRE(ctx && rgb, "tpt_env_set_rgb32f: NULL argument");
    TPT_REQUIRE(w > 0 && h > 0, "tpt_env_set_rgb32f: bad size");
    const size_t n = (size_t)w * h;
    // texels must be finite and in [0, 65536]: one sample's contribution then stays far inside the int64 fixed-point range
    std::vector<float4> staged(n);
    for (size_t i = 0; i < n; ++i) {
        const float r = rgb[3 * i], g = rgb[3 * i + 1], b = rgb[3 * i + 2];
        if (!(r >= 0.0f && r <= 65536.0f && g >= 0.0f && g <= 65536.0f && b >= 0.0f && b <= 65536.0f)) {
            set_error("tpt_env_set_rgb32f: texel " + std::to_string(i) + " is negative, non-finite or above 65536");
            return TPT_ERR_INVALID;
        }
        staged[i] = make_float4(r, g, b, 0.0f);
    }
    TPT_CHECK(bind(ctx));
    TPT_CHECK(env_drop(ctx));
    TPT_CUDA(cudaMalloc(&ctx->envf, n * sizeof(float4)));
    TPT_CUDA(cudaMemcpyAsync(ctx->envf, staged.data(), n * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->env_w = w;
    ctx->env_h = h;
    return TPT_OK;
}

int tpt_env_clear(tpt_context* ctx) {
    TPT_REQUIRE(ctx, "tpt_env_clear: ctx is NULL");
    TPT_CHECK(bind(ctx));
    return env_drop(ctx);
}

int tpt_env_sample(tpt_context* ctx, size_t n, uint64_t seed, float* dirs_pdf) {
    TPT_REQUIRE(ctx && (n == 0 || dirs_pdf), "tpt_env_sample: NULL argument");
    TPT_REQUIRE(ctx->env != nullptr || ctx->envf != nullptr, "tpt_env_sample: no environment map set");
    TPT_CHECK(bind(ctx));
    if (n == 0) return TPT_OK;
    return tpt_env_sample_impl(ctx, n, seed, dirs_pdf);
}

int tpt_env_sampler_tables(tpt_context* ctx, float* prob, uint32_t* alias, float* mass) {
    TPT_REQUIRE(ctx && prob && alias && mass, "tpt_env_sampler_tables: NULL argument");
    TPT_REQUIRE(ctx->env != nullptr || ctx->envf != nullptr, "tpt_env_sampler_tables: no environment map set");
    TPT_CHECK(bind(ctx));
    TPT_CHECK(tpt_env_build_sampler(ctx));
    const size_t n = (size_t)ctx->env_w * ctx->env_h;
    TPT_CUDA(cudaMemcpyAsync(prob, ctx->env_prob, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    TPT_CUDA(cudaMemcpyAsync(alias, ctx->env_alias, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    TPT_CUDA(cudaMemcpyAsync(mass, ctx->env_mass, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    return TPT_OK;
}

static int trace_common(tpt_scene* s, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs, int counted) {
    TPT_REQUIRE(s && (n == 0 || (rays && hits)), "tpt_trace_rays: NULL argument");
    if (!s->built) {
        set_error("tpt_trace_rays: call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    if (n == 0) return TPT_OK;
    tpt_context* ctx = s->ctx;
    TPT_CHECK(bind(ctx));
    if (device_ptrs) return tpt_trace_rays_impl(s, rays, n, hits, any_hit, counted);
    TPT_CHECK(ensure_io(ctx, n));
    // Host buffers: the batch is cut into chunks and pipelined over three streams — chunk i+1 goes up while chunk i is traced
    // and the hits of chunk i-1 come down (PCIe is full duplex; with pinned host memory the call costs about
    // max(copy in + copy out of one direction's worth, traversal) instead of their sum).  Results do not depend on the cut.
    if (!ctx->io_in) {
        TPT_CUDA(cudaStreamCreateWithFlags(&ctx->io_in, cudaStreamNonBlocking));
        TPT_CUDA(cudaStreamCreateWithFlags(&ctx->io_out, cudaStreamNonBlocking));
    }
    const size_t chunk = std::max<size_t>((size_t)1 << 20, (n + 7) / 8);
    const size_t nChunks = (n + chunk - 1) / chunk;
    while (ctx->io_events.size() < 2 * nChunks + 1) {
        cudaEvent_t e;
        TPT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->io_events.push_back(e);
    }
    tpt_ray* dRays = static_cast<tpt_ray*>(ctx->io_rays);
    tpt_hit* dHits = static_cast<tpt_hit*>(ctx->io_hits);
    // the staging buffers may still be in use by earlier work on the context stream
    TPT_CUDA(cudaEventRecord(ctx->io_events[2 * nChunks], ctx->stream));
    TPT_CUDA(cudaStreamWaitEvent(ctx->io_in, ctx->io_events[2 * nChunks], 0));
    for (size_t c = 0; c < nChunks; ++c) {
        const size_t off = c * chunk, cnt = std::min(chunk, n - off);
        TPT_CUDA(cudaMemcpyAsync(dRays + off, rays + off, cnt * sizeof(tpt_ray), cudaMemcpyHostToDevice, ctx->io_in));
        TPT_CUDA(cudaEventRecord(ctx->io_events[2 * c], ctx->io_in));
        TPT_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->io_events[2 * c], 0));
        TPT_CHECK(tpt_trace_rays_impl(s, dRays + off, cnt, dHits + off, any_hit, counted));
        TPT_CUDA(cudaEventRecord(ctx->io_events[2 * c + 1], ctx->stream));
        TPT_CUDA(cudaStreamWaitEvent(ctx->io_out, ctx->io_events[2 * c + 1], 0));
        TPT_CUDA(cudaMemcpyAsync(hits + off, dHits + off, cnt * sizeof(tpt_hit), cudaMemcpyDeviceToHost, ctx->io_out));
    }
    TPT_CUDA(cudaStreamSynchronize(ctx->io_out));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    return TPT_OK;
}

int tpt_trace_rays(tpt_scene* s, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs) {
    return trace_common(s, rays, n, hits, any_hit, device_ptrs, 0);
}
int tpt_trace_rays_counted(tpt_scene* s, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs) {
    return trace_common(s, rays, n, hits, any_hit, device_ptrs, 1);
}

int tpt_generate_primary_rays(tpt_context* ctx, const tpt_camera* cam, int w, int h, int jitter, uint64_t seed, tpt_ray* rays, int device_ptrs) {
    TPT_REQUIRE(ctx && cam && rays, "tpt_generate_primary_rays: NULL argument");
    TPT_REQUIRE(w > 0 && h > 0, "tpt_generate_primary_rays: bad size");
    TPT_CHECK(bind(ctx));
    const size_t n = (size_t)w * h;
    if (device_ptrs) return tpt_primary_rays_impl(ctx, cam, w, h, jitter, seed, rays);
    TPT_CHECK(ensure_io(ctx, n));
    TPT_CHECK(tpt_primary_rays_impl(ctx, cam, w, h, jitter, seed, static_cast<tpt_ray*>(ctx->io_rays)));
    TPT_CUDA(cudaMemcpyAsync(rays, ctx->io_rays, n * sizeof(tpt_ray), cudaMemcpyDeviceToHost, ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(ctx->stream));
    return TPT_OK;
}

int tpt_accum_create(tpt_context* ctx, int w, int h, tpt_accum** out) {
    TPT_REQUIRE(ctx && out, "tpt_accum_create: NULL argument");
    TPT_REQUIRE(w > 0 && h > 0, "tpt_accum_create: bad size");
    TPT_CHECK(bind(ctx));
    tpt_accum* a = new (std::nothrow) tpt_accum();
    if (!a) return TPT_ERR_NOMEM;
    a->ctx = ctx;
    a->width = w;
    a->height = h;
    a->owned = true;
    const size_t bytes = (size_t)w * h * 3 * sizeof(unsigned long long);
    cudaError_t e = cudaMalloc(&a->data, bytes);
    if (e == cudaSuccess) e = cudaMemsetAsync(a->data, 0, bytes, ctx->stream);
    if (e != cudaSuccess) {
        delete a;
        return cuda_fail(e, "accum alloc", __FILE__, __LINE__);
    }
    *out = a;
    return TPT_OK;
}

int tpt_accum_wrap(tpt_context* ctx, void* device_i64, int w, int h, tpt_accum** out) {
    TPT_REQUIRE(ctx && out && device_i64, "tpt_accum_wrap: NULL argument");
    TPT_REQUIRE(w > 0 && h > 0, "tpt_accum_wrap: bad size");
    // the all-reduce kernel reads and writes the buffer with 128-bit accesses (comm.cu), the renderer with 64-bit atomics
    TPT_REQUIRE(reinterpret_cast<uintptr_t>(device_i64) % 16 == 0, "tpt_accum_wrap: the buffer must be 16-byte aligned");
    tpt_accum* a = new (std::nothrow) tpt_accum();
    if (!a) return TPT_ERR_NOMEM;
    a->ctx = ctx;
    a->width = w;
    a->height = h;
    a->owned = false;
    a->data = static_cast<unsigned long long*>(device_i64);
    *out = a;
    return TPT_OK;
}

int tpt_accum_destroy(tpt_accum* a) {
    if (!a) return TPT_OK;
    if (a->owned) {
        cudaSetDevice(a->ctx->device);
        cudaStreamSynchronize(a->ctx->stream);
        cudaFree(a->data);
    }
    delete a;
    return TPT_OK;
}

int tpt_accum_clear(tpt_accum* a) {
    TPT_REQUIRE(a, "tpt_accum_clear: NULL argument");
    TPT_CHECK(bind(a->ctx));
    TPT_CUDA(cudaMemsetAsync(a->data, 0, (size_t)a->width * a->height * 3 * sizeof(unsigned long long), a->ctx->stream));
    return TPT_OK;
}

int tpt_accum_add(tpt_accum* dst, const tpt_accum* src) {
    TPT_REQUIRE(dst && src, "tpt_accum_add: NULL argument");
    TPT_REQUIRE(dst->ctx == src->ctx && dst->width == src->width && dst->height == src->height, "tpt_accum_add: mismatched buffers");
    TPT_CHECK(bind(dst->ctx));
    return tpt_accum_add_impl(dst, src);
}

void* tpt_accum_device_ptr(tpt_accum* a) { return a ? a->data : nullptr; }

int tpt_accum_export_f32(tpt_accum* a, float* host_rgb) {
    TPT_REQUIRE(a && host_rgb, "tpt_accum_export_f32: NULL argument");
    TPT_CHECK(bind(a->ctx));
    const size_t n = (size_t)a->width * a->height * 3;
    float* d = nullptr;
    TPT_CUDA(cudaMalloc(&d, n * sizeof(float)));
    int rc = tpt_accum_to_f32_impl(a, d);
    cudaError_t e = cudaSuccess;
    if (rc == TPT_OK) e = cudaMemcpyAsync(host_rgb, d, n * sizeof(float), cudaMemcpyDeviceToHost, a->ctx->stream);
    if (rc == TPT_OK && e == cudaSuccess) e = cudaStreamSynchronize(a->ctx->stream);
    cudaFree(d);
    TPT_CHECK(rc);
    TPT_CUDA(e);
    return TPT_OK;
}

int tpt_accum_export_i64(tpt_accum* a, int64_t* host) {
    TPT_REQUIRE(a && host, "tpt_accum_export_i64: NULL argument");
    TPT_CHECK(bind(a->ctx));
    TPT_CUDA(cudaMemcpyAsync(host, a->data, (size_t)a->width * a->height * 3 * sizeof(int64_t), cudaMemcpyDeviceToHost, a->ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(a->ctx->stream));
    return TPT_OK;
}

int tpt_accum_import_i64(tpt_accum* a, const int64_t* host) {
    TPT_REQUIRE(a && host, "tpt_accum_import_i64: NULL argument");
    TPT_CHECK(bind(a->ctx));
    TPT_CUDA(cudaMemcpyAsync(a->data, host, (size_t)a->width * a->height * 3 * sizeof(int64_t), cudaMemcpyHostToDevice, a->ctx->stream));
    TPT_CUDA(cudaStreamSynchronize(a->ctx->stream));
    return TPT_OK;
}

int tpt_render(tpt_scene* s, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc) {
    TPT_REQUIRE(s && cam && cfg && acc, "tpt_render: NULL argument");
    TPT_REQUIRE(cfg->width > 0 && cfg->height > 0 && cfg->width <= 16384 && cfg->height <= 16384, "tpt_render: bad resolution");
    TPT_REQUIRE(cfg->spp > 0, "tpt_render: spp must be positive");
    TPT_REQUIRE(cfg->max_depth >= 1 && cfg->max_depth <= 64, "tpt_render: max_depth must be in 1..64");
    TPT_REQUIRE(cfg->sample_stride >= 1 && cfg->sample_offset >= 0, "tpt_render: bad sample slice");
    TPT_REQUIRE(acc->width == cfg->width && acc->height == cfg->height, "tpt_render: accumulator size differs from the render size");
    TPT_REQUIRE(acc->ctx == s->ctx, "tpt_render: accumulator belongs to another context");
    if (!s->built) {
        set_error("tpt_render: call tpt_bvh_build first");
        return TPT_ERR_STATE;
    }
    TPT_CHECK(bind(s->ctx));
    return tpt_render_impl(s, cam, cfg, acc);
}

int tpt_resolve_bgra8(tpt_accum* a, int spp_total, uint8_t* fb, int device_ptr) {
    TPT_REQUIRE(a && fb, "tpt_resolve_bgra8: NULL argument");
    TPT_REQUIRE(spp_total > 0, "tpt_resolve_bgra8: spp_total must be positive");
    TPT_CHECK(bind(a->ctx));
    if (device_ptr) return tpt_resolve_impl(a, spp_total, fb);
    const size_t bytes = (size_t)a->width * a->height * 4;
    uint8_t* d = nullptr;
    TPT_CUDA(cudaMalloc(&d, bytes));
    // the reference leaves the alpha byte as it finds it: start from the caller's buffer
    cudaError_t e = cudaMemcpyAsync(d, fb, bytes, cudaMemcpyHostToDevice, a->ctx->stream);
    int rc = e == cudaSuccess ? tpt_resolve_impl(a, spp_total, d) : TPT_OK;
    if (e == cudaSuccess && rc == TPT_OK) e = cudaMemcpyAsync(fb, d, bytes, cudaMemcpyDeviceToHost, a->ctx->stream);
    if (e == cudaSuccess && rc == TPT_OK) e = cudaStreamSynchronize(a->ctx->stream);
    cudaFree(d);
    TPT_CHECK(rc);
    TPT_CUDA(e);
    return TPT_OK;
}

int tpt_do_trace(tpt_scene* s, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc, uint8_t* host_fb) {
    TPT_REQUIRE(s && cam && cfg && acc && host_fb, "tpt_do_trace: NULL argument");
    if ((cfg->flags & TPT_RENDER_REFIT) && s->topology) TPT_CHECK(tpt_bvh_refit(s, nullptr));
    else TPT_CHECK(tpt_bvh_build(s, nullptr));
    TPT_CHECK(tpt_accum_clear(acc));
    TPT_CHECK(tpt_render(s, cam, cfg, acc));
    return tpt_resolve_bgra8(acc, cfg->spp, host_fb, 0);
}

int tpt_do_trace_device(tpt_scene* s, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc, uint8_t* device_fb) {
    TPT_REQUIRE(s && cam && cfg && acc && device_fb, "tpt_do_trace_device: NULL argument");
    if ((cfg->flags & TPT_RENDER_REFIT) && s->topology) TPT_CHECK(tpt_bvh_refit(s, nullptr));
    else TPT_CHECK(tpt_bvh_build(s, nullptr));
    TPT_CHECK(tpt_accum_clear(acc));
    TPT_CHECK(tpt_render(s, cam, cfg, acc));
    return tpt_resolve_bgra8(acc, cfg->spp, device_fb, 1);
}

int tpt_pointer_is_device(const void* p) {
    cudaPointerAttributes a;
    if (p == nullptr || cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? 1 : 0;
}

}   // extern "C"
