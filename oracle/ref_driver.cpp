// oracle/ref_driver.cpp — TEST INFRASTRUCTURE, not product code.
//
// A small driver around the UNMODIFIED reference sources (compiled where they lie under
// /root/reference/raytracer/src by oracle/Makefile into oracle/_ref/).  It replaces the
// reference's main.cpp (which needs OpenCV C++ headers that this image lacks) and adds
// what the parity tests need and the reference does not expose:
//   * raw float64 framebuffer dump of Tracer::render            (tracer.cpp:247-305)
//   * the primary-hit object-ID buffer, obtained by replaying the very same rays through
//     Scene::intersect (scene.cpp:10-24) with the tmin/tmax of tracer.cpp:29
//   * a band mode (rows [y0,y1) of the Standard loop, tracer.cpp:288-299) that first
//     discards 32*nx*y0 raw MT19937 outputs so a band reproduces the full-frame stream
//   * a "trace these rays" mode for known-answer tests on arbitrary rays.
// Only tests/, bench.py's cpu_baseline / --impl reference leg and __graft_entry__.smoke()
// may execute the binary built from this file.
//
// File formats (little endian):
//   fb   : int32 W, int32 H, then W*H*3 float64 (row 0 = top, as Tracer::render stores)
//   ids  : int32 W, int32 H, int32 S (samples per pixel: 8 std / 1 paper), then W*H*S int32
//          code = obj_index*8 + region (region: 0 none, 1 top, 2 bottom, 3 belt, 4 ring, 5 button)
//          or -1 on miss; followed by W*H*S float64 hit t (inf on miss)
//   rays in : int32 N, then N*6 float64 (ox,oy,oz,dx,dy,dz)
//   rays out: int32 N, then per ray 16 float64:
//          hit, t, px,py,pz, nx,ny,nz, front_face, code, r,g,b (Tracer::trace), occluded(tmin=1e-3,tmax=t_query), 2 pad
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <random>
#include <sstream>
#include <string>
#include <vector>

#include "core.h"
#include "camera.h"
#include "scene.h"
#include "geometry.h"
#include "tracer.h"
#include "json_loader.h"

static int region_code(const Scene& sc, const Hit& h, int obj) {
    int region = 0;
    if (obj >= 0) {
        if (auto pb = dynamic_cast<const Pokeball*>(sc.objects[obj])) {
            if (h.mat == &pb->topMat) region = 1;
            else if (h.mat == &pb->bottomMat) region = 2;
            else if (h.mat == &pb->beltMat) region = 3;
            else if (h.mat == &pb->ringMat) region = 4;
            else if (h.mat == &pb->buttonMat) region = 5;
        }
    }
    return obj * 8 + region;
}

// Same loop as Scene::intersect (scene.cpp:15-21) but remembering the winner's index.
static int closest_index(const Scene& sc, const Ray& r, double tmin, double tmax, Hit& out) {
    Hit temp; int best = -1; double closest = tmax;
    for (size_t i = 0; i < sc.objects.size(); ++i) {
        const Primitive* obj = sc.objects[i];
        if (!obj) continue;
        if (obj->intersect(r, tmin, closest, temp)) { best = (int)i; closest = temp.t; out = temp; }
    }
    return best;
}

static bool write_file(const std::string& path, const std::vector<char>& bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    size_t n = fwrite(bytes.data(), 1, bytes.size(), f);
    fclose(f);
    return n == bytes.size();
}
template <class T> static void put(std::vector<char>& b, const T* p, size_t n) {
    const char* c = reinterpret_cast<const char*>(p);
    b.insert(b.end(), c, c + n * sizeof(T));
}

static void usage() {
    std::cerr <<
      "usage: ref_driver <scene.json> [--paper] [--fb out.fb] [--ids out.ids] [--rows y0 y1] [--stride k off]\n"
      "                  [--rays in.rays out.rays] [--quiet] [--repeat n] [--dirlight dx dy dz r g b]...\n";
}

int main(int argc, char** argv) {
    if (argc < 2) { usage(); return 1; }
    std::string scene_path = argv[1], fb_path, ids_path, rays_in, rays_out;
    bool paper = false, quiet = false; int y0 = -1, y1 = -1, repeat = 1, stride = 1, stride_off = 0;
    std::vector<DirectionalLight> dir_lights;      // the JSON format has none (json_loader.cpp:191-204); pushed into scene.dir_lights below
    for (int i = 2; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--paper") paper = true;
        else if (a == "--quiet") quiet = true;
        else if (a == "--fb" && i + 1 < argc) fb_path = argv[++i];
        else if (a == "--ids" && i + 1 < argc) ids_path = argv[++i];
        else if (a == "--rows" && i + 2 < argc) { y0 = atoi(argv[++i]); y1 = atoi(argv[++i]); }
        else if (a == "--stride" && i + 2 < argc) { stride = atoi(argv[++i]); stride_off = atoi(argv[++i]); if (y0 < 0) { y0 = 0; y1 = 1 << 30; } }
        else if (a == "--rays" && i + 2 < argc) { rays_in = argv[++i]; rays_out = argv[++i]; }
        else if (a == "--repeat" && i + 1 < argc) repeat = atoi(argv[++i]);
        else if (a == "--dirlight" && i + 6 < argc) {
            DirectionalLight L;
            L.dir = Dir3(atof(argv[i + 1]), atof(argv[i + 2]), atof(argv[i + 3]));
            L.radiance = Color{atof(argv[i + 4]), atof(argv[i + 5]), atof(argv[i + 6])};
            dir_lights.push_back(L); i += 6;
        }
        else { usage(); return 1; }
    }
    Scene scene; Camera cam;
    try {
        if (!jsonio::load_scene_from_json(scene_path, scene, cam)) return 2;
    } catch (const std::exception& e) {
        std::cerr << "[error] " << e.what() << "\n";
        return 3;
    }
    for (const auto& L : dir_lights) scene.dir_lights.push_back(L);
    const int W = cam.screen.nx(), H = cam.screen.ny();
    Tracer tracer; tracer.scene = &scene; tracer.camera = &cam; tracer.width = W; tracer.height = H;
    tracer.mode = paper ? RenderMode::Paper : RenderMode::Standard;

    // Tracer::render prints a progress bar to std::cout; silence it if asked.
    std::ostringstream sink; std::streambuf* old = nullptr;
    if (quiet) old = std::cout.rdbuf(sink.rdbuf());

    if (!rays_in.empty()) {
        FILE* f = fopen(rays_in.c_str(), "rb");
        if (!f) { std::cerr << "cannot open rays\n"; return 4; }
        int32_t n = 0; if (fread(&n, 4, 1, f) != 1) return 4;
        std::vector<double> in((size_t)n * 6);
        if (fread(in.data(), 8, in.size(), f) != in.size()) return 4;
        fclose(f);
        std::vector<double> out((size_t)n * 16, 0.0);
        for (int i = 0; i < n; ++i) {
            const double* q = &in[(size_t)i * 6];
            Ray r(Point3(q[0], q[1], q[2]), Dir3(q[3], q[4], q[5]));
            Hit h; int obj = closest_index(scene, r, 1e-4, kINF, h);
            double* o = &out[(size_t)i * 16];
            o[0] = obj >= 0 ? 1.0 : 0.0;
            o[1] = obj >= 0 ? h.t : kINF;
            if (obj >= 0) {
                o[2] = h.p.x; o[3] = h.p.y; o[4] = h.p.z; o[5] = h.n.x; o[6] = h.n.y; o[7] = h.n.z;
                o[8] = h.front_face ? 1.0 : 0.0;
            }
            o[9] = obj >= 0 ? (double)region_code(scene, h, obj) : -1.0;
            Color c = tracer.trace(r);
            o[10] = c.r; o[11] = c.g; o[12] = c.b;
            o[13] = scene.occluded(r, 1e-3, 5.0) ? 1.0 : 0.0;   // fixed probe segment (1e-3, 5]
        }
        std::vector<char> bytes; put(bytes, &n, 1); put(bytes, out.data(), out.size());
        if (!write_file(rays_out, bytes)) return 4;
        if (quiet) std::cout.rdbuf(old);
        return 0;
    }

    std::vector<Color> fb;
    double best_ms = 1e300;
    if (y0 < 0) {
        for (int it = 0; it < repeat; ++it) {
            auto t0 = std::chrono::steady_clock::now();
            tracer.render(fb);
            auto t1 = std::chrono::steady_clock::now();
            best_ms = std::min(best_ms, std::chrono::duration<double, std::milli>(t1 - t0).count());
        }
    } else {
        // Band of the Standard loop (tracer.cpp:288-299): identical statements, rows [y0,y1) only,
        // RNG advanced by discard() to where the full-frame loop would be (2 raw draws per double).
        if (paper) { std::cerr << "--rows is for standard mode\n"; return 1; }
        if (y0 < 0) y0 = 0;
        if (y1 > H) y1 = H;
        fb.assign((size_t)W * H, Color{0, 0, 0});
        // --stride k off: only loop rows y with y % k == off (a uniform sample of the frame); skipped rows are
        // discard()ed from the RNG, and only the time spent inside rendered rows is reported.
        std::mt19937 rng(12345);
        std::uniform_real_distribution<double> uni(-0.5, 0.5);
        rng.discard((unsigned long long)32 * W * y0);
        double ms = 0.0; int rows_done = 0;
        for (int y = y0; y < y1; ++y) {
            if (stride > 1 && (y % stride) != stride_off) { rng.discard((unsigned long long)32 * W); continue; }
            auto t0 = std::chrono::steady_clock::now();
            for (int x = 0; x < W; ++x) {
                Color acc{0, 0, 0};
                for (int s = 0; s < 8; ++s) {
                    double dx = uni(rng), dy = uni(rng);
                    acc += tracer.trace(cam.generate_ray_subpixel(x, y, dx, dy));
                }
                acc = acc * (1.0 / 8.0);
                fb[(size_t)(H - 1 - y) * W + x] = acc;
            }
            auto t1 = std::chrono::steady_clock::now();
            ms += std::chrono::duration<double, std::milli>(t1 - t0).count();
            ++rows_done;
        }
        best_ms = ms;
        std::fprintf(stderr, "rows_rendered %d\n", rows_done);
    }
    if (quiet) std::cout.rdbuf(old);

    if (!fb_path.empty()) {
        std::vector<char> bytes; int32_t hdr[2] = {W, H}; put(bytes, hdr, 2);
        std::vector<double> v((size_t)W * H * 3);
        for (size_t i = 0; i < fb.size(); ++i) { v[3*i] = fb[i].r; v[3*i+1] = fb[i].g; v[3*i+2] = fb[i].b; }
        put(bytes, v.data(), v.size());
        if (!write_file(fb_path, bytes)) return 4;
    }
    if (!ids_path.empty()) {
        const int S = paper ? 1 : 8;
        std::vector<int32_t> ids((size_t)W * H * S, -1);
        std::vector<double> ts((size_t)W * H * S, kINF);
        const int ya = y0 < 0 ? 0 : y0, yb = y0 < 0 ? H : y1;
        if (paper) {
            for (int y = ya; y < yb; ++y) for (int x = 0; x < W; ++x) {
                Hit h; int obj = closest_index(scene, cam.generate_ray(x, y), 1e-4, kINF, h);
                size_t k = (size_t)y * W + x;
                if (obj >= 0) { ids[k] = region_code(scene, h, obj); ts[k] = h.t; }
            }
        } else {
            std::mt19937 rng(12345);
            std::uniform_real_distribution<double> uni(-0.5, 0.5);
            rng.discard((unsigned long long)32 * W * ya);
            for (int y = ya; y < yb; ++y) for (int x = 0; x < W; ++x) for (int s = 0; s < 8; ++s) {
                double dx = uni(rng), dy = uni(rng);
                Hit h; int obj = closest_index(scene, cam.generate_ray_subpixel(x, y, dx, dy), 1e-4, kINF, h);
                size_t k = ((size_t)(H - 1 - y) * W + x) * 8 + s;
                if (obj >= 0) { ids[k] = region_code(scene, h, obj); ts[k] = h.t; }
            }
        }
        std::vector<char> bytes; int32_t hdr[3] = {W, H, S}; put(bytes, hdr, 3);
        put(bytes, ids.data(), ids.size()); put(bytes, ts.data(), ts.size());
        if (!write_file(ids_path, bytes)) return 4;
    }
    std::printf("{\"width\": %d, \"height\": %d, \"mode\": \"%s\", \"rows\": [%d, %d], \"render_ms\": %.3f}\n",
                W, H, paper ? "paper" : "standard", y0 < 0 ? 0 : y0, y0 < 0 ? H : y1, best_ms);
    return 0;
}
