// scene.cuh — device-side data layout of a scene, its BVH and the per-context state.
//
// HBM layout (all arrays 256-B aligned cudaMalloc blocks, sizes for F faces / V vertices):
//   indices      u32[3F]           as uploaded (DeviceScene.indices)
//   obj_verts    float[3V]         object-space positions      obj_norms float[3V]
//   vert_obj     u32[V]            object id of each vertex (built once per scene)
//   face_mtl     u32[F]            material index of each face  (replaces mtlLinearSearch per hit)
//   world_verts  float[3V]         rebuilt by every tpt_bvh_build   world_norms float[3V]
//   nodes        float4[4*(F-1)]   64-B BVH2 nodes, both child boxes in the parent (Aila–Laine packing):
//                                    n0 = (L.lo.x, L.hi.x, L.lo.y, L.hi.y)   n1 = (R.lo.x, R.hi.x, R.lo.y, R.hi.y)
//                                    n2 = (L.lo.z, L.hi.z, R.lo.z, R.hi.z)   n3 = (left, right, -, -) as int bits
//                                  child >= 0: internal node index; child < 0: leaf ~child (sorted triangle index)
//   tri_geom     float4[3F]        per leaf, in Morton order: (v0.xyz, e1.x) (e1.yz, e2.xy) (e2.z, prim, mtl, -)
//   tri_norm     float4[3F]        per leaf: (n0.xyz, n1.x) (n1.yz, n2.xy) (n2.z, -, -, -)  world shading normals
//   mtl_table    float4[2M]        (base.rgb, emission) (eta, metallic, -, -)
//   lights       tpt_delta_light[L]
// Every traversal read is a 128-bit load at a 16-B aligned address.
#pragma once
#include <vector>

#include "common.cuh"

namespace tpt {

struct DeviceCounters {
    unsigned long long rays_primary, rays_extension, rays_shadow, rays_direct;
    unsigned long long node_visits, tri_tests;
    unsigned long long dropped_nonfinite;
    unsigned long long rays_culled;   // camera samples answered by pixel-frustum culling (counted in rays_primary too)
    unsigned long long rays_candidates;   // camera samples answered by testing their pixel's candidate triangles (in rays_primary too)
    unsigned long long queue_overflow;    // vertices whose rays did not fit their queue (must stay 0: the pools hold the worst case)
    unsigned long long rays_queued;       // rays that went through a wavefront queue (written once, read once); the rest were traced where they were born
};

struct BvhView {
    const float4* nodes;
    const float4* nodes4;     // 4-wide compressed nodes (traverse.cuh), or nullptr when the scene has none
    const float4* tri_geom;
    const float4* tri_norm;
    int n_faces;
};

}   // namespace tpt

struct tpt_scene;
// what the live-pixel table of a context was computed for (render.cu: k_tile_live)
struct LiveKey {
    const tpt_scene* scene;
    uint64_t epoch;
    tpt_camera cam;
    int w, h, candidates, measured;
};

struct tpt_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    // environment map (RGBA8, row 0 = bottom)
    uchar4* env = nullptr;
    float4* envf = nullptr;   // or linear float RGB (HDR, tpt_env_set_rgb32f); at most one of env / envf is set
    int env_w = 0, env_h = 0;
    // alias tables over the map's texels (env.cu; built on first use by TPT_RENDER_ENV_NEE / tpt_env_sample, dropped with the map)
    float* env_prob = nullptr;
    uint32_t* env_alias = nullptr;
    float* env_mass = nullptr;
    // counters
    tpt::DeviceCounters* d_counters = nullptr;
    uint64_t samples = 0;
    uint64_t kernel_launches = 0;
    // wavefront queues (grown on demand): ping-pong extension / secondary ray queues, 3 float4 per item
    float4* q_ext[2] = {nullptr, nullptr};
    float4* q_sec[2] = {nullptr, nullptr};
    size_t ext_capacity = 0, sec_capacity = 0;
    size_t pool_budget = 0;                       // bytes the queue pools may take (min(40 GiB, 60 % of free memory)); 0 = not looked up yet
    void* levels = nullptr;                       // QueueLevel[kMaxLevels] (render.cu)
    unsigned wave_grid[4] = {0, 0, 0, 0};          // persistent grid sizes of k_wave<COUNT, NEE> (148 x resident blocks), index COUNT + 2*NEE
    unsigned wave_loop_grid[4] = {0, 0, 0, 0};     // cooperative grid of k_wave_loop<COUNT, NEE> (0 = not queried yet, 1 = no cooperative launch)
    unsigned long long* trace_cursor = nullptr;    // item cursor of k_trace_rays_persistent
    unsigned trace_grid = 0, trace_grid_counted = 0, trace_grid_wide = 0, trace_grid_wide_counted = 0;
    // scratch for trace_rays with host pointers
    void* io_rays = nullptr;
    void* io_hits = nullptr;
    size_t io_capacity = 0;
    // host-pointer tpt_trace_rays: copy-in and copy-out streams + per-chunk events, so that PCIe transfers in both directions
    // overlap the traversal of the neighbouring chunks
    cudaStream_t io_in = nullptr, io_out = nullptr;
    std::vector<cudaEvent_t> io_events;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // live pixels per 8x4 tile of the last (scene geometry, camera, resolution) rendered: bounds what a batch can emit
    uint8_t* d_tile_live = nullptr;
    size_t tile_live_capacity = 0;
    std::vector<uint8_t> tile_live_host;
    LiveKey live_key = {};
    // per-stage timing (TPT_RENDER_STAGE_TIMES)
    std::vector<cudaEvent_t> stage_events;
    double stage_ms[4] = {0, 0, 0, 0};
    uint64_t stage_launches[4] = {0, 0, 0, 0};
};

struct tpt_scene {
    tpt_context* ctx = nullptr;
    size_t n_faces = 0, n_vertices = 0, n_objects = 0, n_materials = 0, n_lights = 0;
    uint32_t* indices = nullptr;
    float* obj_verts = nullptr;
    float* obj_norms = nullptr;
    uint32_t* vert_obj = nullptr;
    uint32_t* face_mtl = nullptr;
    tpt_mtl_interval* objects = nullptr;
    float* vert_trans = nullptr;
    float* normal_trans = nullptr;
    float4* mtl_table = nullptr;
    // page-locked staging of the material table (a pageable source would make cudaMemcpyAsync wait for the stream): a ring of four
    // buffers, each with the event of the last copy out of it, so that a host running a few frames ahead never waits for one
    float4* h_mtl[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t h_mtl_free[4] = {nullptr, nullptr, nullptr, nullptr};
    unsigned h_mtl_next = 0;
    tpt_delta_light* lights = nullptr;
    // per-build
    float* world_verts = nullptr;
    float* world_norms = nullptr;
    unsigned long long* keys[2] = {nullptr, nullptr};
    uint32_t* vals[2] = {nullptr, nullptr};
    void* sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    int* parent_int = nullptr;    // parent of internal node i
    int* parent_leaf = nullptr;   // parent of leaf i
    uint32_t* flags = nullptr;    // refit arrival counters
    float4* merge_slots = nullptr;   // build: 2 x 32 B per split position, what the first child to arrive leaves for its sibling (bvh.cu, k_leaves_build)
    uint32_t* bounds = nullptr;   // 6 ordered-uint encoded floats (min xyz, max xyz)
    float4* nodes = nullptr;
    float4* nodes4 = nullptr;     // allocated when the scene is large enough for the wide traversal to pay (tpt_scene_create)
    float4* tri_geom = nullptr;
    float4* tri_norm = nullptr;
    bool built = false;
    uint64_t geom_epoch = 0;      // bumped by every build / refit: world-space geometry may have changed
    uint64_t norm_epoch = ~0ull;  // geom_epoch the per-leaf normal records (tri_norm) were written for (large scenes write them lazily)
    bool topology = false;        // hierarchy (children, parents, leaf order, range ends) of the last full build is intact: tpt_bvh_refit may reuse it
    bool any_specular = false;    // some material is a mirror or glass (metallic > 0 or eta > 0): k_primary's specular continuation can apply
    bool any_emissive = false;    // some material has emission > 0 (else the reference's extra "direct" rays add nothing and are not traced)
    tpt::BvhView view() const { return tpt::BvhView{nodes, nodes4, tri_geom, tri_norm, (int)n_faces}; }
};

struct tpt_accum {
    tpt_context* ctx = nullptr;
    int width = 0, height = 0;
    unsigned long long* data = nullptr;   // [h][w][3] int64 fixed point, row 0 = bottom
    bool owned = true;
};
