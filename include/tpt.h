/* tpt.h — C ABI of the B200-native path-tracing core (libtpt_b200.so).
 *
 * This is the drop-in boundary for the hot path of Cyruscxy/TinyPathTracer.  The reference has no
 * FFI; its boundary is a C++ class surface.  Every entry point below names the
 * reference interface it replaces (paths relative to the reference repository root).  The C++
 * mirror classes in include/tpt/ (PathTracer, BVH, Scene, DeviceScene, Camera, EnvLight) are thin
 * wrappers over these calls and turn a non-zero status into std::runtime_error, which is the
 * reference's CUDA_CHECK convention (include/intellisense_cuda.h:14-20).
 *
 * Conventions
 *  - plain pointers and sizes only; all structs are POD with the byte layouts of the reference's
 *    upload formats (Material 60 B, DeltaLight 52 B, MtlInterval 8 B, Mat4 64 B column-major,
 *    Ray 24 B, HitStatus 16 B);
 *  - every function returns TPT_OK (0) or a negative tpt_status; tpt_last_error() returns the
 *    message of the last failure on the calling thread;
 *  - one context per GPU; functions on one context are not re-entrant (the reference is single
 *    threaded, default stream, device 0: src/path_tracer.cu:491-554);
 *  - work is issued on the context's stream and is asynchronous unless stated; host-pointer
 *    outputs are valid on return (the call synchronises the stream);
 *  - there is no CPU fallback: with no usable CUDA device every call fails with TPT_ERR_CUDA.
 */
#ifndef TPT_H
#define TPT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TPT_ABI_VERSION 1

typedef enum tpt_status {
    TPT_OK = 0,
    TPT_ERR_INVALID = -1, /* bad argument */
    TPT_ERR_CUDA = -2,    /* CUDA runtime error (message has the cudaError string) */
    TPT_ERR_STATE = -3,   /* call order (e.g. render before bvh build) */
    TPT_ERR_NOMEM = -4
} tpt_status;

typedef struct tpt_context tpt_context;
typedef struct tpt_scene tpt_scene;
typedef struct tpt_accum tpt_accum;

/* include/ray.h:8-16 `Ray` */
typedef struct tpt_ray {
    float origin[3];
    float dir[3];
} tpt_ray;

/* src/path_tracer.cu:19-26 `HitStatus`: prim = -1 and t = FLT_MAX on a miss */
typedef struct tpt_hit {
    int32_t prim;
    float t;
    float u, v;
} tpt_hit;

/* include/material.h:86-120 `Material` (60 bytes; only base_color, emission, eta, metallic are read
 * by the reference's kernels, src/path_tracer.cu:187-225,374-412); unused[2] is `roughness`, read only
 * under TPT_RENDER_ROUGH_METALS */
typedef struct tpt_material {
    float base_color[3];
    float emission;
    float eta;
    float metallic;
    float unused[9];
} tpt_material;

/* include/mesh.cuh:74-78 `MtlInterval`: first face of an object and its material index */
typedef struct tpt_mtl_interval {
    int32_t begin;
    int32_t mtl_idx;
} tpt_mtl_interval;

/* include/delta_light.h:96-130 `DeltaLight` (52 bytes): type 0 point, 1 directional, 2 spot.
 * a = position (point, spot) or direction (directional); b = spot direction. */
typedef struct tpt_delta_light {
    int32_t type;
    float color[3];
    float intensity;
    float a[3];
    float b[3];
    float cos_outer;
    float inv_cone_diff;
} tpt_delta_light;

/* include/mesh.cuh:80-96 `DeviceScene` as filled by Scene::copySceneToDevice (src/mesh.cu:309-397).
 * All pointers are HOST pointers; the arrays are copied to the device by tpt_scene_create.
 * texCoords are part of the reference's upload but are read by no kernel and are not taken. */
typedef struct tpt_scene_desc {
    const uint32_t* indices;   size_t n_indices;    /* 3 per face, rebased to the global vertex array */
    const float* vertices;     /* object space, xyz per vertex */
    const float* normals;      /* object space, xyz per vertex */
    size_t n_vertices;
    const tpt_material* materials;       size_t n_materials;
    const tpt_mtl_interval* objects;     size_t n_objects;
    const float* vert_trans;   /* n_objects x 16, column-major local-to-world */
    const float* normal_trans; /* n_objects x 16, inverse-transpose of the upper 3x3 */
    const tpt_delta_light* lights;       size_t n_lights;
} tpt_scene_desc;

/* include/camera.h:56-65 + Transform::localToWorld (include/transform.h:16-26): exactly the four
 * values that reach the kernel (src/path_tracer.cu:502-506).  vfov is in radians. */
typedef struct tpt_camera {
    float vfov;
    float aspect;
    float near_plane; /* carried for API parity; unused by the reference's kernels too */
    float c2w[16];    /* column-major camera-to-world */
} tpt_camera;

/* Runtime form of the reference's compile-time constants: DEPTH_TRACE (src/path_tracer.cu:17),
 * nSamplesPerPixel (:559), window extent (include/vkEngine.h:24), time() seed (:493-494). */
typedef struct tpt_render_config {
    int32_t width, height;
    int32_t spp;           /* samples per pixel rendered by THIS call */
    int32_t max_depth;     /* DEPTH_TRACE, 1..64 */
    uint64_t seed;
    int32_t sample_offset; /* first sample index; sample j of this call has index offset + j*stride */
    int32_t sample_stride; /* >= 1; slice s of S renders offset = s, stride = S */
    uint32_t flags;        /* TPT_RENDER_* */
    int32_t paths_in_flight; /* 0 = default pool size */
} tpt_render_config;

#define TPT_RENDER_COUNTERS 1u    /* count node/triangle visits per ray kind (slower; for roofline accounting) */
#define TPT_RENDER_STAGE_TIMES 2u /* CUDA-event time of every wavefront kernel, summed per stage (one sync per batch) */
#define TPT_RENDER_NO_PIXEL_CANDIDATES 8u /* check switch: camera samples walk the tree one by one instead of testing their
                                           * pixel's candidate triangles; the accumulation buffer is bit-identical either way */
#define TPT_RENDER_ROUGH_METALS 32u /* metals (metallic > 0, eta = 0) use their `roughness` (Material float 8, loaded by the
                                    * reference at src/mesh.cu:231 but read by none of its kernels): GGX microfacet reflection, visible-normal
                                    * sampling, alpha = roughness^2, Fresnel = baseColor as for the mirror; alpha < 1e-3 is the reference's
                                    * mirror, bit for bit.  Off by default: the reference's metals are perfect mirrors. */
#define TPT_RENDER_NO_SPECULAR_CONTINUATION 64u /* check switch: k_primary queues every ray of the camera vertex (the ray that leaves a mirror /
                                                * glass vertex, the delta-light shadow rays) instead of tracing them itself in registers;
                                                * same accumulator bits, same ray counts */
#define TPT_RENDER_REFIT 16u      /* tpt_do_trace only: refit (tpt_bvh_refit) instead of rebuilding when the scene has a hierarchy */
#define TPT_RENDER_ENV_NEE 4u     /* sample the environment map from its alias table at every diffuse vertex (one extra shadow ray),
                                   * combined with the reference's cosine-sampled misses by the balance heuristic: same expectation as
                                   * the default estimator, lower variance under maps with small bright regions.  Off by default: the
                                   * reference builds the map's pdf/cdf (src/env_light.cu:10-53) but never samples them. */

typedef struct tpt_bvh_stats {
    float build_ms;          /* device time of the whole build, CUDA events on the context stream */
    uint32_t n_faces;
    uint32_t n_nodes;        /* internal nodes of the traversal tree */
    uint32_t max_depth;
    float sah_cost;
    float bounds_min[3], bounds_max[3];
} tpt_bvh_stats;

typedef struct tpt_counters {
    uint64_t rays_primary, rays_extension, rays_shadow, rays_direct;
    uint64_t node_visits, tri_tests;      /* only with TPT_RENDER_COUNTERS / counted trace */
    uint64_t samples;
    uint64_t dropped_nonfinite;           /* contributions that were NaN/Inf and were not accumulated */
    uint64_t kernel_launches;
    uint64_t rays_culled;                 /* camera samples (part of rays_primary) answered without a traversal: their pixel's
                                           * pyramid misses every leaf box of the tree (k_primary's pixel-frustum culling) */
    uint64_t rays_candidates;             /* camera samples (part of rays_primary) answered by testing the 1..8 candidate triangles
                                           * of their pixel's pyramid instead of walking the tree */
    uint64_t queue_overflow;              /* vertices whose rays did not fit their queue and were dropped; 0 by construction (the
                                           * pools hold the worst case), checked by the tests */
    uint64_t rays_queued;                 /* rays (part of extension + shadow + direct) that passed through a wavefront queue: the
                                           * HBM traffic of the bundled scenes (48-B record written once, read once) */
} tpt_counters;

/* Device time per wavefront stage since the last tpt_reset_counters, measured with CUDA events on the
 * context stream around every launch (only renders with TPT_RENDER_STAGE_TIMES contribute). */
typedef struct tpt_stage_times {
    double ms[4];        /* 0 k_primary (generate + closest hit + shade of the camera vertex), 1 k_wave (all later waves:
                          * extension + shadow + direct rays, shading), 2-3 reserved */
    uint64_t launches[4];
} tpt_stage_times;

const char* tpt_last_error(void);
int tpt_abi_version(void);

/* Replaces the implicit device-0 / default-stream state of the reference. */
int tpt_context_create(int device, tpt_context** out);
int tpt_context_destroy(tpt_context* ctx);
int tpt_context_synchronize(tpt_context* ctx);
/* Use an existing CUDA stream (a cudaStream_t cast to void*; e.g. torch.cuda.current_stream().cuda_stream). */
int tpt_context_set_stream(tpt_context* ctx, void* cuda_stream);
int tpt_get_counters(tpt_context* ctx, tpt_counters* out);
int tpt_reset_counters(tpt_context* ctx);
int tpt_get_stage_times(tpt_context* ctx, tpt_stage_times* out);

/* Replaces Scene::copySceneToDevice (src/mesh.cu:309-397): host arrays -> device. */
int tpt_scene_create(tpt_context* ctx, const tpt_scene_desc* desc, tpt_scene** out);
int tpt_scene_destroy(tpt_scene* scene);
/* Copy a new set of host arrays into the device arrays of an existing scene: what Scene::copySceneToDevice
 * (src/mesh.cu:309-397) does per loaded file, without tpt_scene_create's ~24 allocations and tpt_scene_destroy's frees
 * and synchronisation — for callers that re-upload every frame.  Every count must equal the scene's (TPT_ERR_INVALID
 * otherwise); the scene must be rebuilt (tpt_bvh_build) before it is traced; tpt_bvh_refit is refused until then.
 * Asynchronous on the context stream.  Pageable host arrays may be reused on return (the runtime stages them, and may wait for the
 * stream first); page-locked arrays make the call fully asynchronous and must stay unchanged until the stream has passed the copies. */
int tpt_scene_update(tpt_scene* scene, const tpt_scene_desc* desc);
/* Update per-object transforms (n_objects x 16 each, host pointers); the reference re-reads
 * vertTrans/normalTrans every frame (src/path_tracer.cu:536-537). */
int tpt_scene_set_transforms(tpt_scene* scene, const float* vert_trans, const float* normal_trans);

/* Replaces the `transform` kernel (src/path_tracer.cu:239-263, launched :536) and BVH::construct
 * (src/bvh.cu:304-331, called :541-542): object->world vertices/normals, Morton keys, radix sort,
 * hierarchy, refit.  Asynchronous when stats == NULL. */
int tpt_bvh_build(tpt_scene* scene, tpt_bvh_stats* stats);
/* BVH::construct(thrust::device_vector<Vec3>& vertices, thrust::device_vector<uint32_t>& indices) of the reference
 * (include/bvh.cuh:60-68, src/bvh.cu:304-331) for callers that own WORLD-SPACE geometry on the device, as its PathTracer does
 * (m_wVertices, src/path_tracer.cu:515-542): d_world_vertices = n_vertices x 3 floats, d_indices = 3 x n_faces uint32, both
 * DEVICE pointers.  *scene == NULL: a scene of that size is created first (one object, identity transform, one default
 * material, no normals: it can be built, traced and exported; rendering it shades with zero normals) and returned; otherwise
 * the scene must come from an earlier call with the same sizes.  The arrays are copied (device to device) and the tree is
 * built like tpt_bvh_build.  Asynchronous when stats == NULL. */
int tpt_bvh_build_from_world(tpt_context* ctx, const float* d_world_vertices, size_t n_vertices, const uint32_t* d_indices, size_t n_faces,
                             tpt_scene** scene, tpt_bvh_stats* stats);
/* After tpt_scene_set_transforms on an unchanged mesh: keep the hierarchy of the last tpt_bvh_build (leaf order, children,
 * parents) and redo only what depends on the matrices — object->world transform, triangle/normal records, bottom-up boxes.
 * The reference has no such call (it rebuilds per frame, src/path_tracer.cu:540-542).  Hits are those of a
 * rebuild (the boxes are exact again and the closest hit under the tie rule does not depend on the tree); tree quality follows
 * the motion.  TPT_ERR_STATE when the scene was never built.  Asynchronous when stats == NULL. */
int tpt_bvh_refit(tpt_scene* scene, tpt_bvh_stats* stats);
/* Structural export in the reference's layout (include/bvh.cuh:34-58, 36-byte BVHNode, internal nodes
 * [0,F-1) then leaves [F-1,2F-1), child indices global, leaf k holds its face id): n_nodes must be 2F-1. */
int tpt_bvh_export_reference_nodes(tpt_scene* scene, void* out_nodes36, size_t n_nodes);
/* The leaf order of the last tpt_bvh_build: the sorted Morton keys (m_keys of the reference's BVH, include/bvh.cuh:67, after its
 * thrust::sort_by_key, src/bvh.cu:326) and the face id of every sorted leaf.  HOST outputs of n_faces entries; either may be
 * NULL.  The sort is stable: equal keys keep ascending face id. */
int tpt_bvh_export_order(tpt_scene* scene, uint64_t* keys, uint32_t* faces);
/* World-space vertices / normals after the transform step (host output, n_vertices x 3 each; either may be NULL). */
int tpt_scene_export_world(tpt_scene* scene, float* world_vertices, float* world_normals);

/* Replaces EnvLight(const std::string&) + Texture upload (src/env_light.cu:38-54, src/texture.cu:64-171).
 * rgba8: host pointer, w*h*4 bytes, row 0 = BOTTOM scanline (FreeImage order), RGBA.  The lookup is the
 * reference's tex2DLod point/clamp/normalised fetch (src/path_tracer.cu:288-294). */
int tpt_env_set_rgba8(tpt_context* ctx, const uint8_t* rgba8, int width, int height);
/* HDR environment (the reference's Texture is 8-bit only, src/texture.cu:79-82): linear float RGB, host pointer,
 * w*h*3 floats, row 0 = BOTTOM scanline, same lat-long lookup.  Texels must be finite and in [0, 65536] (TPT_ERR_INVALID
 * otherwise).  Replaces a map set by tpt_env_set_rgba8 and vice versa. */
int tpt_env_set_rgb32f(tpt_context* ctx, const float* rgb, int width, int height);
int tpt_env_clear(tpt_context* ctx); /* black environment (the reference's no-env case is UB) */
/* Replaces EnvLightSample / EnvLightPDF (include/env_light.cuh:21-70, no call site in the reference): draws n directions
 * from the map's luminance*sin(theta) distribution with the device alias tables; dirs_pdf is a HOST buffer of n*4 floats
 * (unit direction xyz, density per unit solid angle).  Sample i uses the counter RNG with (seed, i). */
int tpt_env_sample(tpt_context* ctx, size_t n, uint64_t seed, float* dirs_pdf);
/* The alias tables themselves (width*height entries each, row 0 = bottom): acceptance probability, alias texel,
 * probability mass per texel.  HOST buffers. */
int tpt_env_sampler_tables(tpt_context* ctx, float* prob, uint32_t* alias, float* mass);

/* Replaces traverseBVH (src/path_tracer.cu:61-107) as a batch entry: closest hit with t > 2e-4, or
 * (any_hit != 0) prim = index of some hit / -1.  rays/hits are host pointers unless device_ptrs != 0. */
int tpt_trace_rays(tpt_scene* scene, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs);
/* Same, additionally counting node visits and triangle tests into the context counters. */
int tpt_trace_rays_counted(tpt_scene* scene, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel and same results as
 * tpt_trace_rays, but it allocates nothing, on the device or through any allocating library routine: the scratch the kernel
 * needs (its work cursor; for host buffers also the ray and hit staging) is a file-scope `static __device__` array of
 * 256 + TPT_TRACE_STATIC_MAX_RAYS * 40 bytes (about 2.5 MiB) that every GPU which loads the library holds until exit.
 * n > TPT_TRACE_STATIC_MAX_RAYS is refused with TPT_ERR_INVALID.  rays / hits are host pointers unless device_ptrs != 0
 * (device buffers are used in place; host buffers make the call synchronous).  All calls on one GPU share the array: calls
 * that overlap on the GPU (two contexts / streams on the same device) race; calls on one context are ordered by its stream. */
#define TPT_TRACE_STATIC_MAX_RAYS 65536
int tpt_trace_rays_static(tpt_scene* scene, const tpt_ray* rays, size_t n, tpt_hit* hits, int any_hit, int device_ptrs);
/* sampleRays (src/path_tracer.cu:42-59) for every pixel with (u,v) = (0.5,0.5) (jitter == 0) or the
 * sample-0 jitter of `seed`; rays is a device (device_ptrs != 0) or host buffer of width*height tpt_ray,
 * pixel (x,y) at index y*width + x, row 0 = bottom. */
int tpt_generate_primary_rays(tpt_context* ctx, const tpt_camera* cam, int width, int height, int jitter, uint64_t seed,
                              tpt_ray* rays, int device_ptrs);

/* Accumulation buffer: replaces PathTracer::m_radiance (include/path_tracer.h:27; Spectrum per pixel).
 * Stored as 3 x int64 fixed point (2^-36) per pixel so that sums are exact and order-independent. */
int tpt_accum_create(tpt_context* ctx, int width, int height, tpt_accum** out);
/* Wrap caller-owned device memory of width*height*3 int64 (e.g. a torch tensor used for the NCCL sum). */
int tpt_accum_wrap(tpt_context* ctx, void* device_i64, int width, int height, tpt_accum** out);
int tpt_accum_destroy(tpt_accum* acc);
int tpt_accum_clear(tpt_accum* acc);                      /* thrust::fill at src/path_tracer.cu:534 */
/* Multi-GPU from ONE process (the reference drives its single GPU from one host thread).  tpt_comm_init_all creates
 * one context per GPU 0..ngpu-1 and enables peer access between all of them.  tpt_accum_allreduce leaves in EVERY accumulator
 * (one per context, same size) the exact integer sum of all of them: each GPU reduces 1/n of the buffer by reading its peers'
 * memory over NVLink, then fetches the other reduced parts — no library collective; asynchronous, ordered after the work already
 * queued on each context's stream and before whatever is queued next.  With one process PER GPU (torch.distributed) wrap the
 * buffer instead (tpt_accum_wrap / tpt_accum_device_ptr) and use the framework's all-reduce: int64 sums are exact either way. */
int tpt_comm_init_all(int ngpu, tpt_context** out_contexts);
int tpt_accum_allreduce(tpt_accum** per_gpu, int n);
/* Multi-GPU with one process PER GPU (torchrun): the frame exchange.  Rank r of `world` creates an exchange for its
 * accumulator (one made by tpt_accum_create) and publishes TPT_EXCHANGE_HANDLE_BYTES; the host gathers every rank's bytes in rank
 * order (any transport: torch.distributed, MPI, a file) and hands the world*TPT_EXCHANGE_HANDLE_BYTES to tpt_exchange_connect,
 * which maps the peers' accumulators and framebuffers (CUDA IPC; NVLink peer access).  tpt_exchange_resolve then replaces
 * "all-reduce the accumulators, then copyToFB" (src/path_tracer.cu:451-471 on the summed buffer) by ONE fused pass: after a
 * flag handshake in peer memory every rank sums ITS 1/world of the pixels straight out of all accumulators, converts them
 * exactly as tpt_resolve_bgra8 does (alpha byte = the `alpha` given at creation) and stores them into the framebuffer of
 * rank `root` (every rank's if root == -1) — 24 bytes per pixel cross NVLink once and 4 come back.  Asynchronous on the
 * context's stream, to be called by every rank once per frame after its tpt_render calls; when the call's work has completed
 * on a rank, that rank's accumulator may be cleared again and (root) its framebuffer, tpt_exchange_framebuffer — device
 * memory, width*height*4, row 0 = top — holds the frame.  The accumulators themselves are left unsummed.  A rank that waits
 * longer than 20 s (TPT_EXCHANGE_TIMEOUT_MS) for another gives up instead of hanging the GPU: tpt_exchange_status (synchronous)
 * then returns an error.  Destroy the exchange on every rank before any rank destroys its accumulator.  The reference has one
 * GPU and no counterpart of this. */
#define TPT_EXCHANGE_HANDLE_BYTES 168
typedef struct tpt_exchange tpt_exchange;
int tpt_exchange_create(tpt_context* ctx, tpt_accum* acc, int rank, int world, unsigned char alpha, tpt_exchange** out, void* handle_out);
int tpt_exchange_connect(tpt_exchange* x, const void* all_handles);
int tpt_exchange_resolve(tpt_exchange* x, int spp_total, int root);
void* tpt_exchange_framebuffer(tpt_exchange* x);
int tpt_exchange_read_framebuffer(tpt_exchange* x, uint8_t* host_fb);   /* synchronous copy of this rank's framebuffer to the host */
int tpt_exchange_status(tpt_exchange* x);
int tpt_exchange_destroy(tpt_exchange* x);
int tpt_accum_add(tpt_accum* dst, const tpt_accum* src);   /* dst += src (same context) */
void* tpt_accum_device_ptr(tpt_accum* acc);               /* int64[height][width][3], row 0 = bottom */
int tpt_accum_export_f32(tpt_accum* acc, float* host_rgb); /* radiance SUM, float[height][width][3] */
int tpt_accum_export_i64(tpt_accum* acc, int64_t* host_i64);
int tpt_accum_import_i64(tpt_accum* acc, const int64_t* host_i64);

/* Replaces the `trace` megakernel + setupRandSeed (src/path_tracer.cu:34-40,296-435; launched :513,:548).
 * Adds `spp` samples per pixel to `acc`.  Asynchronous. */
int tpt_render(tpt_scene* scene, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc);

/* Replaces copyToFB (src/path_tracer.cu:451-471): B,G,R bytes, alpha byte untouched, row 0 = top.
 * fb: width*height*4 bytes, device pointer if device_ptr != 0, else host. */
int tpt_resolve_bgra8(tpt_accum* acc, int spp_total, uint8_t* fb, int device_ptr);

/* One-call form of PathTracer::doTrace (src/path_tracer.cu:491-554): transform + BVH build + clear +
 * render + resolve into a HOST framebuffer.  Synchronous. */
int tpt_do_trace(tpt_scene* scene, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc, uint8_t* host_fb);

/* PathTracer::doTrace as the reference's displayer calls it (include/path_tracer.h:34; src/vkEngine.cu:244 hands it a DEVICE
 * pointer to W*H*4 bytes): like tpt_do_trace with the framebuffer on the device; asynchronous on the context stream. */
int tpt_do_trace_device(tpt_scene* scene, const tpt_camera* cam, const tpt_render_config* cfg, tpt_accum* acc, uint8_t* device_fb);
/* 1 if p is device (or managed) memory, 0 if it is host memory: lets a host layer without CUDA headers accept either kind of
 * framebuffer behind one signature, as the reference's doTrace does. */
int tpt_pointer_is_device(const void* p);

#ifdef __cplusplus
}
#endif
#endif /* TPT_H */
