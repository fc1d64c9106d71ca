// ref_headless.cu — headless driver for the UNMODIFIED reference sources (test/benchmark infrastructure only).
//
// Built by oracle/ref_build.py from /root/reference (sources are compiled where they lie; only two constants of
// src/path_tracer.cu are made overridable in a generated copy under oracle/_ref/patched/: DEPTH_TRACE (:17) and
// nSamplesPerPixel (:559)).  This file supplies what the reference needs to link and run without a graphics stack:
//   * VkEngine::init/run/draw/cleanup (declared in include/vkEngine.h:17-93): a cudaMalloc'ed W*H*4 framebuffer and a
//     timed frame loop instead of GLFW + Vulkan + external memory (src/vkEngine.cu, not compiled);
//   * main(): PathTracer(env).render(scene) exactly like src/main.cu:6-14, with the file names from the command line;
//   * extern "C" hooks around the reference's own CUDA_CALLABLE functions compiled for the HOST (traverseBVH,
//     rayHitTriangle, rayHitBBox, reflect/refract/shlickFresnel, getMortonCode, DeltaLight::sample, Vec2UV,
//     Transform::localToWorld, Mat4 inverse/transpose, Scene loader) — used to pin the oracle in this container.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

uint32_t g_tpt_ref_width = 0, g_tpt_ref_height = 0;
int g_tpt_ref_spp = 64;
static int g_frames = 3, g_warmup = 1;
static std::string g_out;
static double g_ms_sum = 0.0;
static std::vector<double> g_ms;
int tpt_ref_spp() { return g_tpt_ref_spp; }

#include "path_tracer.cu"   // the generated copy in oracle/_ref/patched (first on the include path)

// ---- headless VkEngine --------------------------------------------------------------------------------------
void VkEngine::init() {
    CUDA_CHECK(cudaMalloc(&m_cudaFramebufferMem, (size_t)m_windowExtent.width * m_windowExtent.height * 4));
    CUDA_CHECK(cudaMemset(m_cudaFramebufferMem, 0, (size_t)m_windowExtent.width * m_windowExtent.height * 4));
    m_isInitialized = true;
}
void VkEngine::draw(std::function<void(unsigned char* framebuffer)> renderJob) {
    renderJob(m_cudaFramebufferMem);            // src/vkEngine.cu:244
    CUDA_CHECK(cudaDeviceSynchronize());        // copyToFB is not synchronised by doTrace (src/path_tracer.cu:553)
    m_numFrames++;
}
void VkEngine::run(std::function<void(unsigned char* framebuffer)> renderJob) {
    for (int i = 0; i < g_warmup; ++i) draw(renderJob);
    for (int i = 0; i < g_frames; ++i) {
        auto t0 = std::chrono::steady_clock::now();
        draw(renderJob);
        auto t1 = std::chrono::steady_clock::now();
        g_ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        g_ms_sum += g_ms.back();
    }
    if (!g_out.empty()) {
        std::vector<unsigned char> fb((size_t)m_windowExtent.width * m_windowExtent.height * 4);
        CUDA_CHECK(cudaMemcpy(fb.data(), m_cudaFramebufferMem, fb.size(), cudaMemcpyDeviceToHost));
        if (FILE* f = fopen(g_out.c_str(), "wb")) {
            fwrite(fb.data(), 1, fb.size(), f);
            fclose(f);
        }
    }
}
void VkEngine::cleanup() {
    if (m_isInitialized) cudaFree(m_cudaFramebufferMem);
    m_isInitialized = false;
}
void VkEngine::render() {}


// ---- BASELINE.json configs[3]: the reference's own BVH::construct (src/bvh.cu:304-331) and traverseBVH
// (src/path_tracer.cu:61-107) on a pre-flattened world-space triangle soup and two ray files --------------------
// File: uint64 {V, F, nA, nB}; float[3V] world vertices; uint32[3F] indices; float[6 nA] rays A; float[6 nB] rays B.
__global__ void tptRefTraceRays(const float* rays6, size_t n, BVHNode* nodes, Vec3* vertices, uint32_t* indices, int nFaces, HitStatus* hits)
{
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	Ray r(Vec3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]), Vec3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]));
	HitStatus st;
	traverseBVH(r, nodes, vertices, indices, nFaces, st);
	hits[i] = st;
}

static int runSoup(const std::string& file, const std::string& out, int builds, int reps)
{
	FILE* f = fopen(file.c_str(), "rb");
	if (!f) { fprintf(stderr, "cannot open %s\n", file.c_str()); return 2; }
	unsigned long long hdr[4];
	if (fread(hdr, 8, 4, f) != 4) return 2;
	const size_t V = hdr[0], F = hdr[1], nA = hdr[2], nB = hdr[3];
	std::vector<Vec3> hv(V);
	std::vector<uint32_t> hi(3 * F);
	std::vector<float> ra(6 * nA), rb(6 * nB);
	if (fread(hv.data(), 12, V, f) != V || fread(hi.data(), 4, 3 * F, f) != 3 * F || fread(ra.data(), 24, nA, f) != nA || fread(rb.data(), 24, nB, f) != nB) return 2;
	fclose(f);
	thrust::device_vector<Vec3> dv(hv.begin(), hv.end());
	thrust::device_vector<uint32_t> di(hi.begin(), hi.end());
	BVH bvh(F);
	std::vector<double> buildMs;
	for (int b = 0; b < builds; ++b) {
		auto t0 = std::chrono::steady_clock::now();
		bvh.construct(dv, di);                       // synchronises internally (src/bvh.cu:325,328,330)
		auto t1 = std::chrono::steady_clock::now();
		buildMs.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
	}
	std::sort(buildMs.begin(), buildMs.end());
	double ms[2] = {0, 0};
	std::vector<HitStatus> hostHits;
	FILE* fo = out.empty() ? nullptr : fopen(out.c_str(), "wb");
	for (int set = 0; set < 2; ++set) {
		const std::vector<float>& r = set ? rb : ra;
		const size_t n = set ? nB : nA;
		thrust::device_vector<float> dr(r.begin(), r.end());
		thrust::device_vector<HitStatus> dh(n);
		cudaEvent_t e0, e1;
		cudaEventCreate(&e0); cudaEventCreate(&e1);
		for (int k = 0; k < reps + 1; ++k) {          // first pass is the warm-up
			if (k == 1) cudaEventRecord(e0);
			tptRefTraceRays KERNEL_DIM((unsigned)((n + 255) / 256), 256) (thrust::raw_pointer_cast(dr.data()), n, thrust::raw_pointer_cast(bvh.m_nodes.data()),
				thrust::raw_pointer_cast(dv.data()), thrust::raw_pointer_cast(di.data()), (int)F, thrust::raw_pointer_cast(dh.data()));
		}
		cudaEventRecord(e1);
		CUDA_CHECK(cudaDeviceSynchronize());
		float t = 0;
		cudaEventElapsedTime(&t, e0, e1);
		ms[set] = t / reps;
		if (fo) {
			hostHits.resize(n);
			thrust::copy(dh.begin(), dh.end(), hostHits.begin());
			fwrite(hostHits.data(), sizeof(HitStatus), n, fo);
		}
	}
	if (fo) fclose(fo);
	printf("{\"faces\": %zu, \"bvh_build_ms\": %.4f, \"bvh_build_ms_min\": %.4f, \"builds\": %d, \"primary_rays\": %zu, \"primary_ms\": %.4f, \"secondary_rays\": %zu, \"secondary_ms\": %.4f, \"build\": \"%s\"}\n",
		F, buildMs[buildMs.size() / 2], buildMs[0], builds, nA, ms[0], nB, ms[1], TPT_REF_BUILD);
	return 0;
}

#ifndef TPT_REF_NO_MAIN
int main(int argc, char** argv) {
    std::string scene, env, soup;
    int soupBuilds = 1, soupReps = 1;
    for (int i = 1; i + 1 < argc; i += 2) {
        std::string k = argv[i], v = argv[i + 1];
        if (k == "--scene") scene = v;
        else if (k == "--soup") soup = v;
        else if (k == "--soup-builds") soupBuilds = atoi(v.c_str());
        else if (k == "--soup-reps") soupReps = atoi(v.c_str());
        else if (k == "--env") env = v;
        else if (k == "--width") g_tpt_ref_width = (uint32_t)atoi(v.c_str());
        else if (k == "--height") g_tpt_ref_height = (uint32_t)atoi(v.c_str());
        else if (k == "--spp") g_tpt_ref_spp = atoi(v.c_str());
        else if (k == "--frames") g_frames = atoi(v.c_str());
        else if (k == "--warmup") g_warmup = atoi(v.c_str());
        else if (k == "--out") g_out = v;
        else if (k == "--depth") { if (atoi(v.c_str()) != DEPTH_TRACE) { fprintf(stderr, "this binary was built with DEPTH_TRACE=%d\n", DEPTH_TRACE); return 3; } }
    }
    if (!soup.empty()) {
        try { return runSoup(soup, g_out, soupBuilds, soupReps); }
        catch (const std::exception& e) { fprintf(stderr, "reference threw: %s\n", e.what()); return 1; }
    }
    if (scene.empty()) { fprintf(stderr, "usage: --scene file.gltf [--env img] [--width W --height H --spp N --frames K --warmup W --out raw]\n"); return 2; }
    try {
        PathTracer pt(env);          // src/main.cu:11
        pt.render(scene);            // src/main.cu:12 (catches and prints std::exception itself, src/path_tracer.cu:575-578)
    } catch (const std::exception& e) {
        fprintf(stderr, "reference threw: %s\n", e.what());
        return 1;
    }
    if (g_ms.empty()) return 1;
    std::vector<double> s = g_ms;
    std::sort(s.begin(), s.end());
    unsigned long long rays = 0;   // counting build only: traverseBVH calls of ALL frames run (warm-up included)
#ifdef TPT_REF_COUNT_RAYS
    cudaMemcpyFromSymbol(&rays, g_tpt_ref_rays, sizeof rays);
#endif
    printf("{\"ms_per_frame\": %.4f, \"ms_median\": %.4f, \"frames\": %d, \"warmup\": %d, \"width\": %u, \"height\": %u, \"spp\": %d, \"depth\": %d, \"rays_all_frames\": %llu, \"build\": \"%s\"}\n",
           g_ms_sum / g_ms.size(), s[s.size() / 2], (int)g_ms.size(), g_warmup, g_tpt_ref_width ? g_tpt_ref_width : 1920u,
           g_tpt_ref_height ? g_tpt_ref_height : 1080u, g_tpt_ref_spp, DEPTH_TRACE, rays, TPT_REF_BUILD);
    return 0;
}
#endif

// ---- host-callable hooks around the reference's own functions (pins for oracle/oracle.cpp) ---------------------
extern "C" {
float ref_frsqrt(float x) { return frsqrt(x); }                                        // include/math/vec.h:44-57
void ref_normalize(const float* v, float* out) { Vec3 r = normalize(Vec3(v[0], v[1], v[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
int ref_ray_hit_bbox(const float* ray6, const float* mn, const float* mx) {            // include/geometry_queries.h:18-46
    Ray r(Vec3(ray6[0], ray6[1], ray6[2]), Vec3(ray6[3], ray6[4], ray6[5]));
    BBox b(Vec3(mn[0], mn[1], mn[2]), Vec3(mx[0], mx[1], mx[2]));
    return rayHitBBox(r, b) ? 1 : 0;
}
int ref_ray_hit_triangle(const float* ray6, const float* tri9, float* out3) {          // include/geometry_queries.h:66-86
    Ray r(Vec3(ray6[0], ray6[1], ray6[2]), Vec3(ray6[3], ray6[4], ray6[5]));
    Real dist = 0;
    Vec2 uv;
    bool ok = rayHitTriangle(r, Vec3(tri9[0], tri9[1], tri9[2]), Vec3(tri9[3], tri9[4], tri9[5]), Vec3(tri9[6], tri9[7], tri9[8]), dist, uv);
    out3[0] = dist; out3[1] = uv.x; out3[2] = uv.y;
    return ok ? 1 : 0;
}
// traverseBVH (src/path_tracer.cu:61-107) over nodes in the reference's own 36-byte layout; hits = HitStatus[n]
void ref_traverse(const void* nodes36, const float* wverts, const uint32_t* indices, int nFaces, const float* rays6, size_t n, void* hits16) {
    static_assert(sizeof(BVHNode) == 36 && sizeof(HitStatus) == 16, "layouts");
    HitStatus* out = reinterpret_cast<HitStatus*>(hits16);
    for (size_t i = 0; i < n; ++i) {
        Ray r(Vec3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]), Vec3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]));
        HitStatus st;
        traverseBVH(r, (BVHNode*)nodes36, (Vec3*)wverts, (uint32_t*)indices, nFaces, st);
        out[i] = st;
    }
}
void ref_reflect(const float* d, const float* n, float* out) { Vec3 r = reflect(Vec3(d[0], d[1], d[2]), Vec3(n[0], n[1], n[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
void ref_refract(const float* d, const float* n, float ior, float* out6) {             // src/path_tracer.cu:143-163
    Real cosI = 0, eta = 0;
    bool tir = false;
    Vec3 r = refract(Vec3(d[0], d[1], d[2]), Vec3(n[0], n[1], n[2]), ior, cosI, eta, tir);
    out6[0] = r.x; out6[1] = r.y; out6[2] = r.z; out6[3] = cosI; out6[4] = eta; out6[5] = tir ? 1.0f : 0.0f;
}
float ref_schlick(float cosI, float eta) { return shlickFresnel(cosI, eta); }          // src/path_tracer.cu:165-173
int ref_mtl_linear_search(int fid, const void* lut, int n) { return mtlLinearSearch(fid, (MtlInterval*)lut, n); }
void ref_light_sample(const void* light52, const float* p, float* out7) {              // include/delta_light.h:108-127
    static_assert(sizeof(DeltaLight) == 52, "DeltaLight layout");
    DeltaLight L;
    memcpy(&L, light52, 52);
    Incoming in = L.sample(Vec3(p[0], p[1], p[2]));
    out7[0] = in.radiance.r; out7[1] = in.radiance.g; out7[2] = in.radiance.b;
    out7[3] = in.direction.x; out7[4] = in.direction.y; out7[5] = in.direction.z; out7[6] = in.distance;
}
void ref_vec2uv(const float* d, float* uv) { Vec2 r = Vec2UV(Vec3(d[0], d[1], d[2])); uv[0] = r.x; uv[1] = r.y; }   // include/env_light.cuh:72-78
void ref_to_uchar(const float* rgb, float spp, unsigned char* out3) {                  // include/material.h:74-81 via operator/ (:71-73)
    Spectrum s(rgb[0], rgb[1], rgb[2]);
    uchar3 c = (s / spp).toUChar();
    out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}
void ref_trs_to_mat4(const float* t, const float* q, const float* s, float* out16) {   // include/transform.h:28-33
    Transform tr(Vec3(t[0], t[1], t[2]), Quat(q[0], q[1], q[2], q[3]), Vec3(s[0], s[1], s[2]));
    Mat4 m = tr.localToWorld();
    memcpy(out16, &m, 64);
}
void ref_normal_matrix(const float* l2w16, float* out16) {                             // the lambda at src/mesh.cu:371-378
    Mat4 l2w;
    memcpy(&l2w, l2w16, 64);
    auto m = Mat4{Vec4{l2w[0][0], l2w[0][1], l2w[0][2], 0.0f}, Vec4{l2w[1][0], l2w[1][1], l2w[1][2], 0.0f},
                  Vec4{l2w[2][0], l2w[2][1], l2w[2][2], 0.0f}, Vec4{0.0f, 0.0f, 0.0f, 1.0f}};
    Mat4 r = m.transpose().inverse();
    memcpy(out16, &r, 64);
}
void ref_mat4_mul_vec4(const float* m16, const float* v4, float* out4) {
    Mat4 m;
    memcpy(&m, m16, 64);
    Vec4 r = m * Vec4(v4[0], v4[1], v4[2], v4[3]);
    out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}
int ref_depth_trace() { return DEPTH_TRACE; }
}   // extern "C"
