// ray_main.cpp — `ray <scene.json> <output.png> [--paper]`, the reference CLI (main.cpp:42-94)
// over the C ABI.  Same argv, same exit codes (1 usage, 2 load failed, 3 exception/"[error]",
// 4 PNG write failed, 0 ok) and the same stdout lines; the render itself runs on the GPU.
// Extras that do not change the 3-argument surface: RAY_STATS=1 prints a JSON line with timings
// and ray counters to stderr; RAY_DEVICE selects the CUDA device.
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

    const char* dev = std::getenv("RAY_DEVICE");
    if ((rc = rt_init(dev ? std::atoi(dev) : -1)) != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }
    rt_scene* scene = nullptr;
    if ((rc = rt_scene_upload(desc, &scene)) != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }

    if (paper) std::cout << "Rendering in paper mode (" << W << "x" << H << ")\n";     // tracer.cpp:259
    else std::cout << "Rendering with 8 spp (" << W << "x" << H << ")\n";               // tracer.cpp:287
    std::vector<uint8_t> rgb((size_t)W * H * 3);
    rt_stats st{};
    rc = rt_render(scene, paper ? RT_MODE_PAPER : RT_MODE_STANDARD, 0, 1, rgb.data(), nullptr, nullptr, &st);
    if (rc != RT_OK) { std::cerr << "[error] " << rt_last_error() << "\n"; return 3; }
    std::cout << "\nRendering complete!\n";                                             // tracer.cpp:304

    if (rt_write_png_rgb8(out_path.c_str(), rgb.data(), W, H) != RT_OK) {
        std::cerr << "Failed to write PNG: " << out_path << "\n";                       // main.cpp:87
        return 4;
    }
    std::cout << "Wrote " << out_path << " (" << W << "x" << H << ")" << (paper ? " (paper mode)" : "") << "\n";   // main.cpp:91-92
    if (std::getenv("RAY_STATS")) {
        const double rays = (double)(st.primary_rays + st.shadow_rays + st.secondary_rays);
        std::fprintf(stderr, "{\"kernel_ms\": %.3f, \"total_ms\": %.3f, \"primary_rays\": %llu, \"shadow_rays\": %llu, "
                             "\"secondary_rays\": %llu, \"prim_tests\": %llu, \"mrays_per_s\": %.1f}\n",
                     st.kernel_ms, st.total_ms, (unsigned long long)st.primary_rays, (unsigned long long)st.shadow_rays,
                     (unsigned long long)st.secondary_rays, (unsigned long long)st.prim_tests, rays / (st.kernel_ms * 1e3));
    }
    rt_scene_free(scene); rt_compiled_free(compiled); rt_shutdown();
    return 0;
}
