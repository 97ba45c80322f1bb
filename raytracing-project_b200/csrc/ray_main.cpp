// ray_main.cpp — `ray <scene.json> <output.png> [--paper]`, the reference CLI (main.cpp:42-94)
// over the C ABI.  Same argv, same exit codes (1 usage, 2 load failed, 3 exception/"[error]",
// 4 PNG write failed, 0 ok) and the same stdout lines; the render itself runs on the GPU.
// Extras that do not change the 3-argument surface: RAY_STATS=1 prints a JSON line with
// timings and ray counters to stderr; RAY_DEVICE selects the CUDA device; RAY_GPUS=N or RAY_GPUS=0,2,5 renders the
// frame on several GPUs of the box from this one process (rt_render_multi: interleaved tiles, every device stores its
// pixels straight into one pinned host frame).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"

int main(int argc, char** argv) {
    if (argc < 3) {
        std::cerr << "Usage: " << argv[0] << " <scene.json> <output.png> [--paper]\n";
        std::cerr << "  --paper: Enable paper rendering mode with crosshatching\n";
        return 1;
    }
    const std::string json_path = argv[1], out_path = argv[2];
    const bool paper = argc > 3 && std::string(argv[3]) == "--paper";

    rt_compiled* compiled = nullptr;
    int rc = rt_compile_json_file(json_path.c_str(), &compiled);
    if (rc != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }
    const rt_scene_desc* desc = rt_compiled_desc(compiled);
    const int W = desc->width, H = desc->height;

    // device list: RAY_GPUS = a count ("8") or a list ("0,2,5"); else the one device RAY_DEVICE names (default: current)
    std::vector<int> gpus;
    if (const char* g = std::getenv("RAY_GPUS")) {
        const std::string spec(g);
        if (spec.find(',') == std::string::npos) { for (int i = 0; i < std::atoi(g); ++i) gpus.push_back(i); }
        else { size_t pos = 0; while (pos <= spec.size()) { const size_t c = std::min(spec.find(',', pos), spec.size()); if (c > pos) gpus.push_back(std::atoi(spec.substr(pos, c - pos).c_str())); pos = c + 1; } }
        if (gpus.empty()) { std::cerr << "[error] RAY_GPUS names no device\n"; return 3; }
    }
    const char* dev = std::getenv("RAY_DEVICE");
    rc = gpus.empty() ? rt_init(dev ? std::atoi(dev) : -1) : rt_init_devices(gpus.data(), (int)gpus.size());
    if (rc != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }
    // the frame: pinned host memory the render kernels store their RGB8 pixels into directly (no device frame, no copy)
    void* frame = nullptr;
    if ((rc = rt_host_frame_alloc((size_t)W * H * 3, nullptr, &frame)) != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }
    uint8_t* rgb = static_cast<uint8_t*>(frame);
    rt_scene* scene = nullptr;
    if (gpus.size() <= 1 && (rc = rt_scene_upload(desc, &scene)) != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }

    if (paper) std::cout << "Rendering in paper mode (" << W << "x" << H << ")\n";     // tracer.cpp:259
    else std::cout << "Rendering with 8 spp (" << W << "x" << H << ")\n";               // tracer.cpp:287
    rt_stats st{};
    const int mode = paper ? RT_MODE_PAPER : RT_MODE_STANDARD;
    rc = gpus.size() > 1 ? rt_render_multi(desc, mode, gpus.data(), (int)gpus.size(), rgb, &st)
                         : rt_render(scene, mode, 0, 1, rgb, nullptr, nullptr, &st);
    if (rc != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }
    std::cout << "\nRendering complete!\n";                                             // tracer.cpp:304

    if (rt_write_png_rgb8(out_path.c_str(), rgb, W, H) != RT_OK) {
        std::cerr << "Failed to write PNG: " << out_path << "\n";                       // main.cpp:87
        return 4;
    }
    std::cout << "Wrote " << out_path << " (" << W << "x" << H << ")" << (paper ? " (paper mode)" : "") << "\n";   // main.cpp:91-92
    if (std::getenv("RAY_STATS")) {
        const double rays = (double)(st.primary_rays + st.shadow_rays + st.secondary_rays);
        std::fprintf(stderr, "{\"kernel_ms\": %.3f, \"total_ms\": %.3f, \"primary_rays\": %llu, \"shadow_rays\": %llu, "
                             "\"secondary_rays\": %llu, \"prim_tests\": %llu, \"mrays_per_s\": %.1f, \"gpus\": %d}\n",
                     st.kernel_ms, st.total_ms, (unsigned long long)st.primary_rays, (unsigned long long)st.shadow_rays,
                     (unsigned long long)st.secondary_rays, (unsigned long long)st.prim_tests, rays / (st.kernel_ms * 1e3),
                     (int)std::max<size_t>(gpus.size(), 1));
    }
    rt_host_frame_free(frame); rt_scene_free(scene); rt_compiled_free(compiled); rt_shutdown();
    return 0;
}
