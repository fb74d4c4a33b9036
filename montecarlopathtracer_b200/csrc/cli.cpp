// mcpt_cli -- headless driver over the C ABI: the call sequence of the reference's two front ends
// (CVMCTracer/CVMCTracer/main.cpp:9-50: readObj -> CreateGeometry -> RenderScene -> PNG;
//  GraphicsRTX::DoOnUpdate: progressive accumulation with a temp.png every 100 frames) without
// Win32 / D3D11 / OpenCV.
//
//   mcpt_cli --scene FILE.obj|FILE.mcpt [--loader tinyobj|objreader] [--tree sah|reference]
//            [--w W --h H --spp N --depth D] [--eye x,y,z --at x,y,z --fovy DEG] [--emission S]
//            [--twin SCENEID (before --eye/--depth overrides)] [--seed K --rng philox|reference] [--nee] [--step N] [--out out.png] [--pfm out.pfm] [--device I]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mcpt.h"

namespace {
int die(const char* what) {
    std::fprintf(stderr, "mcpt_cli: %s: %s\n", what, mcpt_last_error());
    return 1;
}
bool ends_with(const std::string& s, const char* suf) {
    const size_t n = std::strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}
bool parse3(const char* s, float out[3]) { return std::sscanf(s, "%f,%f,%f", out, out + 1, out + 2) == 3; }
}  // namespace

int main(int argc, char** argv) {
    std::string scene, out = "result.png", pfm;
    int loader = MCPT_LOADER_TINYOBJ, build = MCPT_BUILD_SAH, device = 0, step = 0;
    mcpt_render_cfg cfg;
    mcpt_default_render_cfg(&cfg, 640, 480, 16, 5);  // QuinEngine window (Main.cpp:11), depth literal rtx.hlsl:400
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "mcpt_cli: %s needs a value\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--scene") scene = val();
        else if (a == "--loader") loader = std::string(val()) == "objreader" ? MCPT_LOADER_OBJREADER : MCPT_LOADER_TINYOBJ;
        else if (a == "--tree") build = std::string(val()) == "reference" ? MCPT_BUILD_REFERENCE : MCPT_BUILD_SAH;
        else if (a == "--w") cfg.width = std::atoi(val());
        else if (a == "--h") cfg.height = std::atoi(val());
        else if (a == "--spp") cfg.spp = std::atoi(val());
        else if (a == "--depth") cfg.max_depth = std::atoi(val());
        else if (a == "--eye") { if (!parse3(val(), cfg.eye)) return 2; }
        else if (a == "--at") { if (!parse3(val(), cfg.at)) return 2; }
        else if (a == "--fovy") cfg.fovy = static_cast<float>(std::atof(val()) * 3.14159265358979323846 / 180.0);
        else if (a == "--emission") cfg.emission_scale = static_cast<float>(std::atof(val()));
        else if (a == "--seed") cfg.seed = static_cast<uint32_t>(std::strtoul(val(), nullptr, 10));
        else if (a == "--rng") cfg.rng_mode = std::string(val()) == "reference" ? MCPT_RNG_REFERENCE : MCPT_RNG_PHILOX;
        else if (a == "--nee") cfg.flags |= MCPT_FLAG_NEE;
        else if (a == "--twin") {  // the CUDA tracer's estimator and camera (RenderScene(sceneID, ...))
            const int w = cfg.width, h = cfg.height, n = cfg.spp;
            mcpt_cuda_twin_render_cfg(&cfg, w, h, n, std::atoi(val()));
        }
        else if (a == "--step") step = std::atoi(val());
        else if (a == "--out") out = val();
        else if (a == "--pfm") pfm = val();
        else if (a == "--device") device = std::atoi(val());
        else { std::fprintf(stderr, "mcpt_cli: unknown option %s\n", a.c_str()); return 2; }
    }
    if (scene.empty()) { std::fprintf(stderr, "usage: mcpt_cli --scene FILE.obj|FILE.mcpt [options]  (see csrc/cli.cpp)\n"); return 2; }

    if (mcpt_initialize(device)) return die("initialize");
    const size_t slash = scene.find_last_of('/');
    const std::string base = slash == std::string::npos ? std::string() : scene.substr(0, slash + 1);
    if (ends_with(scene, ".mcpt") ? mcpt_load_scene_file(scene.c_str(), build) : mcpt_load_obj(scene.c_str(), base.c_str(), loader, build))
        return die("load scene");
    mcpt_tree_info ti;
    mcpt_get_tree_info(&ti);
    std::fprintf(stderr, "scene %s: %d nodes, %d leaves, %d leaf refs, depth %d, %d emitting triangles, build %.3f s, %.2f MB on device\n",
                 scene.c_str(), ti.num_nodes, ti.num_leaves, ti.num_leaf_refs, ti.max_depth, ti.num_emitter_triangles, ti.build_seconds,
                 ti.device_bytes / 1048576.0);

    // same limits as mcpt_render_scene's check_cfg, applied BEFORE anything is sized from the command line
    if (cfg.width <= 0 || cfg.height <= 0 || cfg.width > 32768 || cfg.height > 32768 || cfg.spp <= 0) {
        std::fprintf(stderr, "mcpt_cli: --w/--h must be in 1..32768 and --spp positive\n");
        return 2;
    }
    const size_t pixels = static_cast<size_t>(cfg.width) * cfg.height;
    std::vector<float> img(3 * pixels, 0.0f), part(3 * pixels);
    const int total = cfg.spp;
    if (step <= 0 || step > total) step = total;
    double ms = 0, rays = 0;
    int done = 0;
    for (int k = 0; done < total; ++k) {   // progressive: running mean over sample ranges, a step image after each
        const int n = std::min(step, total - done);
        mcpt_render_cfg c = cfg;
        c.spp = n;
        c.sample_offset = cfg.sample_offset + done;
        if (mcpt_render_scene(&c, part.data(), nullptr)) return die("render");
        mcpt_counters cnt;
        mcpt_get_counters(&cnt);
        ms += cnt.total_ms; rays += static_cast<double>(cnt.rays);
        const float wOld = static_cast<float>(done) / (done + n), wNew = static_cast<float>(n) / (done + n);
        for (size_t i = 0; i < 3 * pixels; ++i) img[i] = img[i] * wOld + part[i] * wNew;
        done += n;
        if (step < total) {
            char name[64];
            std::snprintf(name, sizeof name, "step%06d.png", k);   // CUTracer.cu:394
            if (mcpt_write_png(name, img.data(), cfg.width, cfg.height, 2.2f)) return die("write step image");
        }
    }
    if (mcpt_write_png(out.c_str(), img.data(), cfg.width, cfg.height, 2.2f)) return die("write png");
    if (!pfm.empty() && mcpt_write_pfm(pfm.c_str(), img.data(), cfg.width, cfg.height)) return die("write pfm");
    const double paths = static_cast<double>(pixels) * total;
    std::printf("{\"paths\": %.0f, \"rays\": %.0f, \"device_ms\": %.3f, \"mpaths_per_s\": %.2f, \"mrays_per_s\": %.2f, \"out\": \"%s\"}\n",
                paths, rays, ms, paths / ms / 1e3, rays / ms / 1e3, out.c_str());
    mcpt_shutdown();
    return 0;
}
