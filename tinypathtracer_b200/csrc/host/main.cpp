// main.cpp — tpt_render: the reference's main() (src/main.cu:6-14: PathTracer(env).render(gltf)) with its hard-coded
// file names, resolution (include/vkEngine.h:24), spp (src/path_tracer.cu:559) and depth (:17) as arguments.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "tpt/host.hpp"

int main(int argc, char** argv) {
    std::string scene, env, out, outFloat;
    int w = 1920, h = 1080, spp = 64, depth = 8, frames = 1;
    unsigned long long seed = 1;
    bool envNee = false, fullGltf = false, progressive = false, roughMetals = false, refit = false;
    std::string checkpoint, resume, devices;
    int gpus = 1;
    for (int i = 1; i < argc; ++i) {
        const std::string k = argv[i];
        auto val = [&]() -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", k.c_str()); std::exit(2); }
            return argv[++i];
        };
        if (k == "--scene") scene = val();
        else if (k == "--env") env = val();
        else if (k == "--out") out = val();
        else if (k == "--out-float") outFloat = val();
        else if (k == "--w" || k == "--width") w = std::atoi(val());
        else if (k == "--h" || k == "--height") h = std::atoi(val());
        else if (k == "--spp") spp = std::atoi(val());
        else if (k == "--depth") depth = std::atoi(val());
        else if (k == "--frames") frames = std::atoi(val());
        else if (k == "--seed") seed = std::strtoull(val(), nullptr, 10);
        else if (k == "--env-nee") envNee = true;
        else if (k == "--gpus") gpus = std::atoi(val());
        else if (k == "--devices") devices = val();   // e.g. 0,1,2,3 (an id may repeat)
        else if (k == "--rough-metals") roughMetals = true;
        else if (k == "--refit") refit = true;
        else if (k == "--full-gltf") fullGltf = true;
        else if (k == "--progressive") progressive = true;
        else if (k == "--checkpoint") checkpoint = val();
        else if (k == "--resume") resume = val();
        else { std::fprintf(stderr, "unknown argument %s\n", k.c_str()); return 2; }
    }
    if (scene.empty()) {
        std::fprintf(stderr, "usage: tpt_render --scene file.gltf [--env sky.png|sky.jpg|sky.hdr|sky.pfm] [--w 1920 --h 1080 --spp 64 --depth 8 --frames 1 --seed 1 --gpus N | --devices 0,1,.. --env-nee --rough-metals --refit --full-gltf --progressive --checkpoint file --resume file --out frame.png --out-float frame.pfm]\n");
        return 2;
    }
    try {
        tpt::PathTracer pathTracer(env);   // src/main.cu:11
        pathTracer.m_width = w;
        pathTracer.m_height = h;
        pathTracer.m_spp = spp;
        pathTracer.m_depth = depth;
        pathTracer.m_frames = frames;
        pathTracer.m_seed = seed;
        pathTracer.m_envNEE = envNee;
        if (!devices.empty()) {
            for (size_t pos = 0; pos <= devices.size();) {
                const size_t c = devices.find(',', pos);
                pathTracer.m_devices.push_back(std::atoi(devices.substr(pos, c == std::string::npos ? std::string::npos : c - pos).c_str()));
                if (c == std::string::npos) break;
                pos = c + 1;
            }
        } else {
            for (int g = 0; g < gpus && gpus > 1; ++g) pathTracer.m_devices.push_back(g);
        }
        pathTracer.m_roughMetals = roughMetals;
        pathTracer.m_refit = refit;
        if (fullGltf) pathTracer.m_sceneType = "gltf-full";
        pathTracer.m_progressive = progressive;
        if (!resume.empty()) pathTracer.loadCheckpoint(resume);   // continues its sample sequence and seed
        pathTracer.m_output = out;
        pathTracer.m_outputFloat = outFloat;
        pathTracer.render(scene);          // src/main.cu:12
        if (!checkpoint.empty()) pathTracer.saveCheckpoint(checkpoint);
        const tpt_counters c = pathTracer.context().counters();
        std::printf("{\"ms_last_frame\": %.3f, \"samples\": %llu, \"rays\": %llu, \"kernel_launches\": %llu}\n", pathTracer.m_lastFrameMs,
                    (unsigned long long)c.samples, (unsigned long long)(c.rays_primary + c.rays_extension + c.rays_shadow + c.rays_direct),
                    (unsigned long long)c.kernel_launches);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());   // the reference prints e.what() (src/path_tracer.cu:575-578)
        return 1;
    }
    return 0;
}
