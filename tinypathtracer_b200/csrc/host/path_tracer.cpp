// path_tracer.cpp — EnvLight and PathTracer of the reference (include/env_light.cuh:9-19, include/path_tracer.h:16-35,
// src/path_tracer.cu:477-579) over the C ABI, plus the extern "C" hooks the tests use to read the loader's output.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>

#include "tpt/host.hpp"

namespace tpt {

// EnvLight(file): decode, swizzle to RGBA8 and upload with row 0 = the BOTTOM scanline — what FreeImage's bottom-up
// bitmap + the BGR->RGBA transform give the reference's texture (include/picture.h:19-45, src/texture.cu:33-47,64-171).
EnvLight::EnvLight(Context& ctx, const std::string& file) {
    if (file.empty()) {
        check(tpt_env_clear(ctx.handle()), "tpt_env_clear");
        return;
    }
    if (PictureF::isFloatFile(file)) {   // HDR map: linear float texels, rows flipped to the reference's bottom-up order
        const PictureF pic(file);
        std::vector<float> rgb((size_t)pic.width * pic.height * 3);
        for (int y = 0; y < pic.height; ++y) {
            const float* src = &pic.pixels[(size_t)(pic.height - 1 - y) * pic.width * 3];
            float* dst = &rgb[(size_t)y * pic.width * 3];
            // the C ABI takes finite texels in [0, 65536]; files may hold negative or larger values (or NaN)
            for (int i = 0; i < 3 * pic.width; ++i) dst[i] = src[i] >= 0.0f ? std::min(src[i], 65536.0f) : 0.0f;
        }
        check(tpt_env_set_rgb32f(ctx.handle(), rgb.data(), pic.width, pic.height), "tpt_env_set_rgb32f");
        return;
    }
    const Picture pic(file);
    std::vector<uint8_t> rgba((size_t)pic.width * pic.height * 4);
    for (int y = 0; y < pic.height; ++y) {
        const uint8_t* src = &pic.pixels[(size_t)(pic.height - 1 - y) * pic.width * pic.channels];
        uint8_t* dst = &rgba[(size_t)y * pic.width * 4];
        for (int x = 0; x < pic.width; ++x) {
            dst[4 * x + 0] = src[pic.channels * x + 0];
            dst[4 * x + 1] = src[pic.channels * x + 1];
            dst[4 * x + 2] = src[pic.channels * x + 2];
            dst[4 * x + 3] = pic.channels == 4 ? src[4 * x + 3] : 255;
        }
    }
    check(tpt_env_set_rgba8(ctx.handle(), rgba.data(), pic.width, pic.height), "tpt_env_set_rgba8");
}

PathTracer::PathTracer() : PathTracer(std::string()) {}
PathTracer::PathTracer(const std::string& envLightFile) : m_ctx(new Context(0)), m_envFile(envLightFile) {
    EnvLight env(*m_ctx, m_envFile);
}
PathTracer::~PathTracer() {
    if (m_radiance) tpt_accum_destroy(m_radiance);
}

void PathTracer::ensureAccumulator() {
    if (!m_radiance || m_accW != m_width || m_accH != m_height) {
        if (m_radiance) tpt_accum_destroy(m_radiance);
        m_radiance = nullptr;
        check(tpt_accum_create(m_ctx->handle(), m_width, m_height, &m_radiance), "tpt_accum_create");
        m_accW = m_width;
        m_accH = m_height;
        m_samplesDone = 0;
    }
}

namespace {
struct CheckpointHeader {
    char magic[8];   // "TPTACC1\0"
    int32_t width, height;
    uint64_t samples, seed;
};
}   // namespace

void PathTracer::saveCheckpoint(const std::string& file) {
    if (!m_radiance) throw std::runtime_error("saveCheckpoint: nothing rendered yet");
    std::vector<int64_t> acc((size_t)m_accW * m_accH * 3);
    check(tpt_accum_export_i64(m_radiance, acc.data()), "tpt_accum_export_i64");
    CheckpointHeader h;
    std::memset(&h, 0, sizeof h);
    std::memcpy(h.magic, "TPTACC1", 8);
    h.width = m_accW;
    h.height = m_accH;
    h.samples = m_samplesDone;
    h.seed = m_seed;
    FILE* f = std::fopen(file.c_str(), "wb");
    if (!f) throw std::runtime_error("saveCheckpoint: cannot create " + file);
    const bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(acc.data(), sizeof(int64_t), acc.size(), f) == acc.size();
    std::fclose(f);
    if (!ok) throw std::runtime_error("saveCheckpoint: short write to " + file);
}

void PathTracer::loadCheckpoint(const std::string& file) {
    FILE* f = std::fopen(file.c_str(), "rb");
    if (!f) throw std::runtime_error("loadCheckpoint: cannot open " + file);
    CheckpointHeader h;
    if (std::fread(&h, sizeof h, 1, f) != 1 || std::memcmp(h.magic, "TPTACC1", 8) != 0) {
        std::fclose(f);
        throw std::runtime_error("loadCheckpoint: " + file + " is not a checkpoint");
    }
    if (h.width != m_width || h.height != m_height) {
        std::fclose(f);
        throw std::runtime_error("loadCheckpoint: checkpoint is " + std::to_string(h.width) + "x" + std::to_string(h.height) + ", renderer is " +
                                 std::to_string(m_width) + "x" + std::to_string(m_height));
    }
    std::vector<int64_t> acc((size_t)h.width * h.height * 3);
    const bool ok = std::fread(acc.data(), sizeof(int64_t), acc.size(), f) == acc.size();
    std::fclose(f);
    if (!ok) throw std::runtime_error("loadCheckpoint: truncated checkpoint " + file);
    ensureAccumulator();
    check(tpt_accum_import_i64(m_radiance, acc.data()), "tpt_accum_import_i64");
    m_samplesDone = h.samples;
    m_seed = h.seed;
    m_progressive = true;
}

// One frame exactly as src/path_tracer.cu:491-554: transform + BVH rebuild, clear, trace, copyToFB.  Progressive mode
// skips the clear, continues the sample sequence and resolves the mean of everything accumulated.
void PathTracer::doTrace(DeviceScene& scene, Camera& camera, unsigned char* framebuffer, int nSamplesPerPixel) {
    ensureAccumulator();
    tpt_render_config cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.width = m_width;
    cfg.height = m_height;
    cfg.spp = nSamplesPerPixel;
    cfg.max_depth = m_depth;
    cfg.seed = m_seed;
    cfg.sample_offset = 0;
    cfg.sample_stride = 1;
    cfg.flags = (m_envNEE ? TPT_RENDER_ENV_NEE : 0u) | (m_roughMetals ? TPT_RENDER_ROUGH_METALS : 0u) | (m_refit ? TPT_RENDER_REFIT : 0u);
    const tpt_camera cam = camera.toABI();
    const auto t0 = std::chrono::steady_clock::now();
    const int onDevice = tpt_pointer_is_device(framebuffer);
    if (!m_progressive) {
        if (onDevice) check(tpt_do_trace_device(scene.handle.get(), &cam, &cfg, m_radiance, framebuffer), "tpt_do_trace_device");
        else check(tpt_do_trace(scene.handle.get(), &cam, &cfg, m_radiance, framebuffer), "tpt_do_trace");
        m_samplesDone = (uint64_t)nSamplesPerPixel;
    } else {
        cfg.sample_offset = (uint32_t)m_samplesDone;
        check(tpt_bvh_build(scene.handle.get(), nullptr), "tpt_bvh_build");
        check(tpt_render(scene.handle.get(), &cam, &cfg, m_radiance), "tpt_render");
        m_samplesDone += (uint64_t)nSamplesPerPixel;
        check(tpt_resolve_bgra8(m_radiance, (int)m_samplesDone, framebuffer, onDevice), "tpt_resolve_bgra8");
    }
    m_lastFrameMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

void PathTracer::writeOutputs(const std::vector<unsigned char>& fb, tpt_accum* acc, uint64_t samples) {
    if (!m_output.empty()) {
        std::vector<uint8_t> rgb((size_t)m_width * m_height * 3);
        for (size_t i = 0; i < (size_t)m_width * m_height; ++i) {   // framebuffer is B,G,R,(A untouched), row 0 = top
            rgb[3 * i + 0] = fb[4 * i + 2];
            rgb[3 * i + 1] = fb[4 * i + 1];
            rgb[3 * i + 2] = fb[4 * i + 0];
        }
        writePNG(m_output, rgb.data(), m_width, m_height, 3);
    }
    if (!m_outputFloat.empty()) {
        std::vector<float> sum((size_t)m_width * m_height * 3), img(sum.size());
        check(tpt_accum_export_f32(acc, sum.data()), "tpt_accum_export_f32");   // radiance SUM, row 0 = bottom
        const float inv = 1.0f / (float)(samples ? samples : 1);
        for (int y = 0; y < m_height; ++y)
            for (size_t i = 0; i < (size_t)m_width * 3; ++i)
                img[(size_t)y * m_width * 3 + i] = sum[(size_t)(m_height - 1 - y) * m_width * 3 + i] * inv;
        writePFM(m_outputFloat, img.data(), m_width, m_height);
    }
}

void PathTracer::render(const std::string& meshFile) {
    if (m_devices.size() > 1) {
        renderMulti(meshFile);
        return;
    }
    Scene scene(meshFile, m_sceneType);                       // src/path_tracer.cu:558 ("gltf")
    DeviceScene d_scene = scene.copySceneToDevice(*m_ctx);    // :560
    Camera& camera = scene.m_camera;                          // :562
    std::vector<unsigned char> fb((size_t)m_width * m_height * 4, 0);
    // the reference hands this job to VkEngine::run, which calls it once per presented frame until the window closes
    // (src/vkEngine.cu:294-302); headless: m_frames frames, each an independent nSamplesPerPixel render with a new seed
    for (int f = 0; f < m_frames; ++f) {
        doTrace(d_scene, camera, fb.data(), m_spp);
        std::printf("1 frame: %.3f ms\n", m_lastFrameMs);     // src/vkEngine.cu:300 prints "1 frame"
        if (!m_progressive) ++m_seed;
    }
    writeOutputs(fb, m_radiance, m_samplesDone);
}

// The same job on several GPUs driven by this one thread: every call below only queues work on its GPU's stream, so the GPUs
// run concurrently; the exchange is one tpt_accum_allreduce per frame.
void PathTracer::renderMulti(const std::string& meshFile) {
    const int n = (int)m_devices.size();
    if (m_progressive) throw std::runtime_error("PathTracer: progressive mode runs on one GPU");
    if (m_spp % n) throw std::runtime_error("PathTracer: spp must be a multiple of the number of GPUs");
    Scene scene(meshFile, m_sceneType);
    std::vector<std::unique_ptr<Context>> own;   // contexts beyond the constructor's (GPU 0); destroyed after scenes and accumulators
    std::vector<Context*> ctx(n);
    for (int g = 0; g < n; ++g) {
        if (g == 0 && m_devices[0] == 0) {
            ctx[g] = m_ctx.get();
        } else {
            own.emplace_back(new Context(m_devices[g]));
            ctx[g] = own.back().get();
            EnvLight env(*ctx[g], m_envFile);
        }
    }
    struct Accums {
        std::vector<tpt_accum*> a;
        ~Accums() { for (tpt_accum* p : a) if (p) tpt_accum_destroy(p); }
    } acc;
    acc.a.assign(n, nullptr);
    std::vector<DeviceScene> scenes;
    for (int g = 0; g < n; ++g) {
        scenes.push_back(scene.copySceneToDevice(*ctx[g]));
        check(tpt_accum_create(ctx[g]->handle(), m_width, m_height, &acc.a[g]), "tpt_accum_create");
    }
    const tpt_camera cam = scene.m_camera.toABI();
    std::vector<unsigned char> fb((size_t)m_width * m_height * 4, 0);
    for (int f = 0; f < m_frames; ++f) {
        const auto t0 = std::chrono::steady_clock::now();
        for (int g = 0; g < n; ++g) {
            tpt_render_config cfg;
            std::memset(&cfg, 0, sizeof cfg);
            cfg.width = m_width;
            cfg.height = m_height;
            cfg.spp = m_spp / n;
            cfg.max_depth = m_depth;
            cfg.seed = m_seed;
            cfg.sample_offset = g;
            cfg.sample_stride = n;
            cfg.flags = (m_envNEE ? TPT_RENDER_ENV_NEE : 0u) | (m_roughMetals ? TPT_RENDER_ROUGH_METALS : 0u);
            tpt_scene* s = scenes[g].handle.get();
            if (m_refit && f > 0) check(tpt_bvh_refit(s, nullptr), "tpt_bvh_refit");
            else check(tpt_bvh_build(s, nullptr), "tpt_bvh_build");
            check(tpt_accum_clear(acc.a[g]), "tpt_accum_clear");
            check(tpt_render(s, &cam, &cfg, acc.a[g]), "tpt_render");
        }
        check(tpt_accum_allreduce(acc.a.data(), n), "tpt_accum_allreduce");
        check(tpt_resolve_bgra8(acc.a[0], m_spp, fb.data(), 0), "tpt_resolve_bgra8");
        m_lastFrameMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::printf("1 frame: %.3f ms\n", m_lastFrameMs);
        ++m_seed;
    }
    m_samplesDone = (uint64_t)m_spp;
    for (int g = 1; g < n; ++g) ctx[g]->synchronize();
    writeOutputs(fb, acc.a[0], m_samplesDone);
    scenes.clear();   // before the contexts they live in
}

}   // namespace tpt

// ---- C hooks for the tests: the loader's flattened arrays -----------------------------------------------------------
extern "C" {
struct tpt_host_scene {
    tpt::DeviceScene d;
    tpt::Camera cam;
    tpt_camera abi;
    std::string error;
};
static thread_local std::string g_host_error;
const char* tpt_host_last_error(void) { return g_host_error.c_str(); }

tpt_host_scene* tpt_host_load_gltf(const char* path) {
    try {
        tpt::Scene scene(path, "gltf");
        tpt_host_scene* h = new tpt_host_scene();
        h->d = scene.flatten();
        h->cam = scene.m_camera;
        h->abi = h->cam.toABI();
        return h;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return nullptr;
    }
}
tpt_host_scene* tpt_host_load_scene(const char* path, const char* type) {   // type: "gltf" | "gltf-full"
    try {
        tpt::Scene scene(path, type);
        tpt_host_scene* h = new tpt_host_scene();
        h->d = scene.flatten();
        h->cam = scene.m_camera;
        h->abi = h->cam.toABI();
        return h;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return nullptr;
    }
}
void tpt_host_scene_free(tpt_host_scene* h) { delete h; }
// sizes: {n_indices, n_vertices, n_materials, n_objects, n_lights}
void tpt_host_scene_sizes(const tpt_host_scene* h, size_t* out5) {
    out5[0] = h->d.indices.size();
    out5[1] = h->d.vertices.size();
    out5[2] = h->d.materials.size();
    out5[3] = h->d.materialsLUT.size();
    out5[4] = h->d.lights.size();
}
void tpt_host_scene_copy(const tpt_host_scene* h, uint32_t* indices, float* vertices, float* normals, float* texcoords, void* materials,
                         void* objects, float* vertTrans, float* normalTrans, void* lights, tpt_camera* cam) {
    const tpt::DeviceScene& d = h->d;
    std::memcpy(indices, d.indices.data(), d.indices.size() * 4);
    std::memcpy(vertices, d.vertices.data(), d.vertices.size() * 12);
    std::memcpy(normals, d.normals.data(), d.normals.size() * 12);
    std::memcpy(texcoords, d.texCoords.data(), d.texCoords.size() * 8);
    std::memcpy(materials, d.materials.data(), d.materials.size() * sizeof(tpt_material));
    std::memcpy(objects, d.materialsLUT.data(), d.materialsLUT.size() * sizeof(tpt_mtl_interval));
    std::memcpy(vertTrans, d.vertTrans.data(), d.vertTrans.size() * 64);
    std::memcpy(normalTrans, d.normalTrans.data(), d.normalTrans.size() * 64);
    if (!d.lights.empty()) std::memcpy(lights, d.lights.data(), d.lights.size() * sizeof(tpt_delta_light));
    *cam = h->abi;
}
// PNG round trip through the repository's codecs (width*height*channels bytes, top-down rows)
int tpt_host_write_png(const char* path, const uint8_t* px, int w, int h, int channels) {
    try {
        tpt::writePNG(path, px, w, h, channels);
        return 0;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -1;
    }
}
int tpt_host_read_image_f32(const char* path, float* out, size_t capacityFloats, int* wh2) {
    try {
        const tpt::PictureF pic(path);
        wh2[0] = pic.width;
        wh2[1] = pic.height;
        if (out) {
            if (capacityFloats < pic.pixels.size()) throw std::runtime_error("tpt_host_read_image_f32: buffer too small");
            std::memcpy(out, pic.pixels.data(), pic.pixels.size() * sizeof(float));
        }
        return 0;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -1;
    }
}

int tpt_host_write_pfm(const char* path, const float* rgbTopDown, int w, int h) {
    try {
        tpt::writePFM(path, rgbTopDown, w, h);
        return 0;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -1;
    }
}

int tpt_host_read_image(const char* path, uint8_t* out, size_t capacity, int* whc3) {
    try {
        tpt::Picture p(path);
        whc3[0] = p.width; whc3[1] = p.height; whc3[2] = p.channels;
        if (out && p.pixels.size() <= capacity) std::memcpy(out, p.pixels.data(), p.pixels.size());
        return 0;
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -1;
    }
}
}   // extern "C"
