// host.hpp — C++ mirrors of the reference's host classes over the C ABI of include/tpt.h.
//
// Same names, argument meaning and error behaviour as the reference (a non-zero status becomes
// std::runtime_error, its CUDA_CHECK convention, include/intellisense_cuda.h:14-20):
//   Transform   include/transform.h:8-59        Camera      include/camera.h:9-65
//   Mesh        include/mesh.cuh:16-70          Scene       include/mesh.cuh:98-115, src/mesh.cu:68-397
//   DeviceScene include/mesh.cuh:80-96          BVH         include/bvh.cuh:60-68, src/bvh.cu:304-331
//   EnvLight    include/env_light.cuh:9-19      PathTracer  include/path_tracer.h:16-35, src/path_tracer.cu:477-579
//   Picture     include/picture.h:12-55 (FreeImage in the reference; PNG/PPM codecs of this repository here)
// No CUDA header is needed to use these: device work happens behind libtpt_b200.so.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../tpt.h"

namespace tpt {

using Real = float;
struct Vec2 { Real x = 0, y = 0; };
struct Vec3 {
    Real x = 0, y = 0, z = 0;
    Vec3() = default;
    explicit Vec3(Real s) : x(s), y(s), z(s) {}
    Vec3(Real a, Real b, Real c) : x(a), y(b), z(c) {}
};
struct Quat {   // include/math/quat.h: (x, y, z, w); the default (all zero) rotates like the identity
    Real x = 0, y = 0, z = 0, w = 0;
    Quat() = default;
    Quat(Real a, Real b, Real c, Real d) : x(a), y(b), z(c), w(d) {}
};
struct Mat4 {   // column-major, include/math/mat.h:37-51
    Real m[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    Real& at(int col, int row) { return m[col * 4 + row]; }
    Real at(int col, int row) const { return m[col * 4 + row]; }
};
Mat4 operator*(const Mat4& a, const Mat4& b);

using Material = tpt_material;        // include/material.h:86-120, 60 bytes
using DeltaLight = tpt_delta_light;   // include/delta_light.h:96-130, 52 bytes
using MtlInterval = tpt_mtl_interval; // include/mesh.cuh:74-78
Material defaultMaterial();           // Material() of include/material.h:88-103

class Transform {
public:
    Transform() = default;
    Transform(Vec3 loc, Quat quat, Vec3 s) : m_location(loc), m_rotation(quat), m_scale(s) {}
    Mat4 localToParent() const;   // Translate * Rotate * Scale, include/transform.h:28-33 (or the node's `matrix`, "gltf-full" only)
    // The reference declares m_parent (include/transform.h:55) but never sets it (src/mesh.cu:265 TODO), so with its loader
    // ("gltf") this is localToParent(); the "gltf-full" loader links the node hierarchy.
    Mat4 localToWorld() const { return m_parent ? m_parent->localToWorld() * localToParent() : localToParent(); }
    Vec3 m_location{0, 0, 0};
    Quat m_rotation;
    Vec3 m_scale{1, 1, 1};
    std::shared_ptr<Transform> m_parent;
    bool m_hasMatrix = false;     // glTF `matrix` property given instead of TRS
    Mat4 m_matrix;
};
Mat4 normalToWorld(const Mat4& localToWorld);   // inverse transpose of the upper 3x3, src/mesh.cu:371-378

class Camera {
public:
    Camera() : m_transform(std::make_shared<Transform>()) {}
    Camera(Real fov, Real asp, Real nearZ, const Transform& trans)
        : m_transform(std::make_shared<Transform>(trans)), m_verticalFOV(fov), m_aspectRatio(asp), m_nearPlane(nearZ) {}
    Real getVFov() const { return m_verticalFOV; }        // radians, straight from glTF yfov (src/mesh.cu:148)
    Real getAspRatio() const { return m_aspectRatio; }
    Real getNearPlane() const { return m_nearPlane; }
    tpt_camera toABI() const;
    std::shared_ptr<Transform> m_transform;

private:
    Real m_verticalFOV = 60.0f, m_aspectRatio = 1.77778f, m_nearPlane = 0.1f;
};

struct Mesh {
    std::vector<uint32_t> m_polygons;
    std::vector<Vec3> m_vertexCoords, m_vertexNormals;
    std::vector<Vec2> m_vertexTexCoords;
    std::string m_material;
    std::shared_ptr<Transform> m_transform;
};

[[noreturn]] void throwLastError(const char* where);
inline void check(int status, const char* where) {
    if (status != TPT_OK) throwLastError(where);
}

// One per GPU: replaces the reference's implicit device 0 / default stream.
class Context {
public:
    explicit Context(int device = 0);
    ~Context();
    // one process, several GPUs: contexts for GPUs 0..n-1 with peer access between all of them (tpt_comm_init_all);
    // sum per-GPU accumulators with tpt_accum_allreduce
    static std::vector<std::unique_ptr<Context>> initAll(int nGpus);
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    tpt_context* handle() const { return m_ctx; }
    void synchronize();
    tpt_counters counters();
    // The reference has no context: device 0 and the default stream are implicit (src/path_tracer.cu:491-554).  Its signatures
    // without a context argument (Scene::copySceneToDevice(), EnvLight(file), BVH(nFaces)) use this process-wide one on device 0.
    static Context& defaultContext();

private:
    explicit Context(tpt_context* adopted) : m_ctx(adopted) {}
    tpt_context* m_ctx = nullptr;
};

// The flattened arrays of the reference's DeviceScene; the device copy lives behind `handle`.
struct DeviceScene {
    std::vector<uint32_t> indices;
    std::vector<Vec3> vertices, normals;
    std::vector<Vec2> texCoords;
    std::vector<Material> materials;
    std::vector<MtlInterval> materialsLUT;
    std::vector<Mat4> vertTrans, normalTrans;
    std::vector<DeltaLight> lights;
    std::shared_ptr<tpt_scene> handle;   // tpt_scene_create / tpt_scene_destroy
    size_t nFaces() const { return indices.size() / 3; }
    // push the current vertTrans / normalTrans to the device (the reference re-reads them every frame, src/path_tracer.cu:536-537);
    // follow with BVH::construct() or BVH::refit()
    void uploadTransforms();
};

class Scene {
public:
    Scene() = default;
    // type "gltf": the reference's loader, quirks included (src/mesh.cu:68-307: flat node list, TRS only, primitives[0] only,
    // materials and lights keyed by name).  type "gltf-full" (SURVEY §8(f)-1): the scene graph of `scenes[scene]` with parent
    // transforms and `matrix` nodes, every triangle primitive of a mesh, instanced meshes, bufferView.byteStride,
    // non-indexed primitives, computed normals where NORMAL is absent, materials and lights keyed by index.
    // Both accept the .glb container as well as .gltf + external / base64 buffers.
    Scene(const std::string& filename, const std::string& type);
    DeviceScene flatten() const;                                   // host part of copySceneToDevice (src/mesh.cu:309-397)
    DeviceScene copySceneToDevice(Context& ctx) const;             // + tpt_scene_create
    DeviceScene copySceneToDevice() const { return copySceneToDevice(Context::defaultContext()); }   // the reference's signature (include/mesh.cuh:103)
    Camera m_camera;
    const std::vector<Mesh>& meshes() const { return m_meshes; }
    const std::map<std::string, Material>& materials() const { return m_materials; }
    const std::map<std::string, DeltaLight>& lights() const { return m_lights; }

private:
    void readFromGLTF(const std::string& filename, bool full);
    std::vector<Mesh> m_meshes;
    std::map<std::string, Material> m_materials;   // std::map: name order, as uploaded by the reference (src/mesh.cu:326-333)
    std::map<std::string, DeltaLight> m_lights;
};

struct BVHNode36 {   // include/bvh.cuh:34-58, the reference's 36-byte node (structural export only)
    uint32_t parent;
    int32_t a, b;
    float mn[3], mx[3];
};

class BVH {
public:
    explicit BVH(DeviceScene& scene) : m_scene(scene.handle) {}
    // The reference's own form (include/bvh.cuh:60-68): BVH bvh(nFaces); bvh.construct(m_wVertices, d_scene.indices) with
    // WORLD-SPACE vertices and indices in thrust::device_vectors (src/path_tracer.cu:541-542).  Any pair of containers whose
    // data() yields device pointers works (thrust::raw_pointer_cast is found by argument-dependent lookup; this header needs no
    // thrust); elements: 12-byte vertices, uint32 indices.
    explicit BVH(size_t nFaces) : m_nFaces(nFaces) {}
    template <class DeviceVertices, class DeviceIndices>
    void construct(DeviceVertices& vertices, DeviceIndices& indices) {
        static_assert(sizeof(*raw_pointer_cast(vertices.data())) == 12 && sizeof(*raw_pointer_cast(indices.data())) == 4, "Vec3 / uint32_t elements");
        constructFromWorld(reinterpret_cast<const float*>(raw_pointer_cast(vertices.data())), vertices.size(),
                           reinterpret_cast<const uint32_t*>(raw_pointer_cast(indices.data())), indices.size() / 3);
    }
    void constructFromWorld(const float* deviceWorldVertices, size_t nVertices, const uint32_t* deviceIndices, size_t nFaces,
                            Context& ctx = Context::defaultContext());   // tpt_bvh_build_from_world
    void construct();                       // `transform` kernel + BVH::construct of the reference
    tpt_bvh_stats constructWithStats();
    void refit();                           // after new object matrices on the same mesh: keeps the hierarchy
    tpt_bvh_stats refitWithStats();
    std::vector<BVHNode36> exportReferenceLayout(size_t nFaces);
    void traceRays(const tpt_ray* rays, size_t n, tpt_hit* hits, bool anyHit = false);   // host buffers

private:
    std::shared_ptr<tpt_scene> m_scene;
    size_t m_nFaces = 0;
};

// 8-bit image, top-down rows, 3 or 4 channels (include/picture.h loads through FreeImage; here: PNG, baseline JPEG, binary PPM).
struct Picture {
    int width = 0, height = 0, channels = 0;
    std::vector<uint8_t> pixels;
    Picture() = default;
    explicit Picture(const std::string& filename);
};
void writePNG(const std::string& filename, const uint8_t* rgba, int width, int height, int channels);

// Float image, top-down rows, linear RGB (HDR environment maps and float dumps of the accumulation buffer; the
// reference's Picture/Texture are 8-bit only).  Reads Radiance RGBE (.hdr, flat or run-length coded scanlines) and PFM.
struct PictureF {
    int width = 0, height = 0;
    std::vector<float> pixels;   // width * height * 3
    PictureF() = default;
    explicit PictureF(const std::string& filename);
    static bool isFloatFile(const std::string& filename);   // by magic: "#?" (Radiance) or "PF"/"Pf"
};
void writePFM(const std::string& filename, const float* rgbTopDown, int width, int height);   // little-endian "PF"

class EnvLight {
public:
    EnvLight(Context& ctx, const std::string& file);   // "" -> black environment (the reference: uninitialised texture)
    explicit EnvLight(const std::string& file) : EnvLight(Context::defaultContext(), file) {}   // the reference's signature (include/env_light.cuh:12)
};

class PathTracer {
public:
    PathTracer();
    explicit PathTracer(const std::string& envLightFile);
    void render(const std::string& meshFile);   // load, upload, m_frames x doTrace, PNG of the last frame if m_output is set
    // framebuffer: W*H*4 bytes, B,G,R,(A untouched), row 0 = top — a DEVICE pointer as the reference's displayer passes it
    // (include/path_tracer.h:34, src/vkEngine.cu:244; the call is then asynchronous like the reference's copyToFB launch) or a host
    // pointer (filled on return); told apart at run time (tpt_pointer_is_device)
    void doTrace(DeviceScene& scene, Camera& camera, unsigned char* framebuffer, int nSamplesPerPixel);

    // run-time forms of the reference's compile-time constants (include/vkEngine.h:24, src/path_tracer.cu:17,559,493)
    int m_width = 1920, m_height = 1080;
    int m_spp = 64, m_depth = 8, m_frames = 1;
    uint64_t m_seed = 1;
    // Progressive mode (the reference restarts every frame, src/path_tracer.cu:534): frames ADD their samples —
    // frame f renders sample indices [f*spp, (f+1)*spp) with the same seed, and the framebuffer shows the mean of all samples so
    // far.  Accumulation is exact (int64), so N progressive frames equal one frame of N*spp samples bit for bit.
    bool m_progressive = false;
    uint64_t m_samplesDone = 0;          // samples per pixel in the accumulator (progressive mode / after loadCheckpoint)
    // Checkpoint = the int64 accumulation buffer + {width, height, samples done, seed}; resuming continues the sample sequence
    // (SURVEY §8(f)-2), so an interrupted render finishes with the image an uninterrupted one gives.
    void saveCheckpoint(const std::string& file);
    void loadCheckpoint(const std::string& file);   // switches to progressive mode
    std::string m_sceneType = "gltf";   // "gltf" = the reference's loader; "gltf-full" = scene graph, all primitives (Scene)
    bool m_roughMetals = false;   // TPT_RENDER_ROUGH_METALS: metals scatter by GGX from Material::roughness (the reference: mirrors)
    bool m_refit = false;         // TPT_RENDER_REFIT: frames after the first keep the BVH hierarchy and refit it (the reference rebuilds)
    // Several GPUs from this one host thread: device ids, one context each (an id may repeat: several contexts on one
    // GPU, which is how the single-GPU test runs it).  Rank g renders the sample indices = g (mod n); the accumulators are summed
    // by tpt_accum_allreduce (peer-memory kernel) and GPU m_devices[0] resolves: the frame is bit-identical for any n.
    std::vector<int> m_devices;
    bool m_envNEE = false;   // TPT_RENDER_ENV_NEE: sample the environment map at diffuse vertices (same expectation, lower variance)
    std::string m_output;
    std::string m_outputFloat;   // PFM of the linear mean radiance (accumulator / samples), unclamped; empty = none
    double m_lastFrameMs = 0.0;
    Context& context() { return *m_ctx; }

private:
    void renderMulti(const std::string& meshFile);
    void writeOutputs(const std::vector<unsigned char>& fb, tpt_accum* acc, uint64_t samples);
    void ensureAccumulator();
    std::unique_ptr<Context> m_ctx;
    std::string m_envFile;
    tpt_accum* m_radiance = nullptr;
    int m_accW = 0, m_accH = 0;

public:
    ~PathTracer();
};

}   // namespace tpt
