// rt_host_api.cpp — host-only half of the C ABI: error channel, scene compiler entry points
// (the replacement for jsonio::load_scene_from_json, json_loader.cpp:458-504) and the PNG writer
// (the replacement for cv::imwrite, main.cpp:84-89; zlib only, no OpenCV).
#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "rt_internal.hpp"
#include "mt19937.cuh"

namespace { thread_local std::string g_error; }
void rt_set_error(const std::string& message) { g_error = message; }

// definition lives in scene_compiler.cpp
struct rt_compiled;
namespace rtb { void destroy_compiled(rt_compiled*); const rt_scene_desc* desc_of(const rt_compiled*); void add_dir_light(rt_compiled*, const DirectionalLight&); }

extern "C" {

const char* rt_last_error(void) { return g_error.c_str(); }

int rt_compile_json_text(const char* json_text, rt_compiled** out) {
    if (!json_text || !out) { rt_set_error("rt_compile_json_text: null argument"); return RT_ERR_ARG; }
    *out = nullptr;
    try {
        rtb::Scene scene; rtb::Camera cam;
        jsonio::load_scene_from_json_text(json_text, scene, cam);
        try { *out = rtb::compile_scene(scene, cam); }
        catch (const std::exception& e) { rt_set_error(e.what()); return RT_ERR_LIMIT; }
        return RT_OK;
    } catch (const std::exception& e) { rt_set_error(e.what()); return RT_ERR_PARSE; }
}

int rt_compile_json_file(const char* path, rt_compiled** out) {
    if (!path || !out) { rt_set_error("rt_compile_json_file: null argument"); return RT_ERR_ARG; }
    *out = nullptr;
    try {
        rtb::Scene scene; rtb::Camera cam;
        jsonio::load_scene_from_json(path, scene, cam);
        try { *out = rtb::compile_scene(scene, cam); }
        catch (const std::exception& e) { rt_set_error(e.what()); return RT_ERR_LIMIT; }
        return RT_OK;
    } catch (const std::exception& e) {
        rt_set_error(e.what());
        return std::strncmp(e.what(), "Cannot open", 11) == 0 ? RT_ERR_IO : RT_ERR_PARSE;
    }
}

const rt_scene_desc* rt_compiled_desc(const rt_compiled* c) { return c ? rtb::desc_of(c) : nullptr; }
void rt_compiled_free(rt_compiled* c) { if (c) rtb::destroy_compiled(c); }

int rt_compiled_add_dir_light(rt_compiled* c, const double dir[3], const double radiance[3]) {
    if (!c || !dir || !radiance) { rt_set_error("rt_compiled_add_dir_light: null argument"); return RT_ERR_ARG; }
    rtb::DirectionalLight L;
    L.dir = {dir[0], dir[1], dir[2]}; L.radiance = {radiance[0], radiance[1], radiance[2]};
    rtb::add_dir_light(c, L);
    return RT_OK;
}

int rt_debug_mt_block_host(uint32_t seed, uint64_t block, uint32_t* out624) {
    if (!out624) { rt_set_error("rt_debug_mt_block_host: null argument"); return RT_ERR_ARG; }
    uint32_t a[mt::kN], b[mt::kN];
    mt::seed_state(seed, a);
    uint32_t *cur = a, *nxt = b;
    // even blocks by the word-parallel form (next_word: the checkpoint / jump kernels), odd ones by the thread form of the
    // render kernels' tile prologue (twist_by_thread): both must reproduce std::mt19937
    for (uint64_t i = 0; i <= block; ++i) {
        if (i & 1) for (int t = 0; t < 227; ++t) mt::twist_by_thread(cur, t, [nxt](int k, uint32_t w) { nxt[k] = w; });
        else for (int k = 0; k < mt::kN; ++k) nxt[k] = mt::next_word(cur, k);
        uint32_t* t = cur; cur = nxt; nxt = t;
    }
    for (int k = 0; k < mt::kN; ++k) out624[k] = mt::temper(cur[k]);
    return RT_OK;
}

int rt_debug_mt_jump_host(uint32_t seed, int log2_twists, uint32_t* out624) {
    if (!out624 || log2_twists < 0 || log2_twists >= mt::kJumpPolys) { rt_set_error("rt_debug_mt_jump_host: bad argument"); return RT_ERR_ARG; }
    std::vector<uint32_t> x((size_t)mt::kJumpBlocks * mt::kN);
    mt::seed_state(seed, x.data());
    for (int b = 0; b + 1 < mt::kJumpBlocks; ++b)
        for (int k = 0; k < mt::kN; ++k) x[(size_t)(b + 1) * mt::kN + k] = mt::next_word(x.data() + (size_t)b * mt::kN, k);
    uint32_t st[mt::kN];
    for (int w = 0; w < mt::kN; ++w) st[w] = mt::jump_word(x.data(), mt::kJumpPoly[log2_twists], w);
    for (int k = 0; k < mt::kN; ++k) out624[k] = mt::temper(mt::next_word(st, k));
    return RT_OK;
}

// 8-bit RGB PNG, no alpha (main.cpp:19-34,84-89 writes CV_8UC3 through cv::imwrite), filter 0 on every row, one zlib
// stream in one IDAT chunk.  The stream is deflated in parallel: bands of rows are compressed independently as raw deflate
// (each ends on a full flush, the last one finishes the stream) on the host's cores, concatenated behind one zlib header,
// and the Adler-32 of the whole is combined from the bands' — a 3840x2160 frame takes ~30 ms on 8 cores instead of ~430 ms,
// which was most of the wall time of `ray scene.json out.png` once the frame itself renders in 12 ms.  Level 1, OpenCV's
// default for PNG; the decoded pixels do not depend on it.
namespace {
struct PngBand { std::vector<unsigned char> z; uLong adler{1}; size_t raw_bytes{0}; bool ok{true}; };

void deflate_band(const uint8_t* rgb8, size_t stride, int y0, int y1, bool last, PngBand* out) {
    std::vector<unsigned char> raw((stride + 1) * (size_t)(y1 - y0));
    for (int y = y0; y < y1; ++y) {
        unsigned char* row = &raw[(stride + 1) * (size_t)(y - y0)];
        row[0] = 0;                                                    // filter type None
        std::memcpy(row + 1, rgb8 + stride * (size_t)y, stride);
    }
    out->raw_bytes = raw.size();
    out->adler = adler32(adler32(0L, Z_NULL, 0), raw.data(), (uInt)raw.size());
    z_stream zs; std::memset(&zs, 0, sizeof zs);
    if (deflateInit2(&zs, 1, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { out->ok = false; return; }
    out->z.resize(deflateBound(&zs, (uLong)raw.size()) + 16);
    zs.next_in = raw.data(); zs.avail_in = (uInt)raw.size();
    zs.next_out = out->z.data(); zs.avail_out = (uInt)out->z.size();
    const int rc = deflate(&zs, last ? Z_FINISH : Z_FULL_FLUSH);
    out->ok = last ? rc == Z_STREAM_END : (rc == Z_OK && zs.avail_in == 0 && zs.avail_out != 0);
    out->z.resize(out->z.size() - zs.avail_out);
    deflateEnd(&zs);
}
}  // namespace

int rt_write_png_rgb8(const char* path, const uint8_t* rgb8, int width, int height) {
    if (!path || !rgb8 || width <= 0 || height <= 0) { rt_set_error("rt_write_png_rgb8: bad argument"); return RT_ERR_ARG; }
    const size_t stride = (size_t)width * 3;
    // bands of >= 64 rows and >= 256 KB, at most one per core; every band's raw size must fit zlib's 32-bit counters
    const unsigned cores = std::max(1u, std::thread::hardware_concurrency());
    const size_t total = (stride + 1) * (size_t)height;
    int n_bands = (int)std::min<size_t>({(size_t)cores, (size_t)(height + 63) / 64, total / (256u << 10) + 1});
    n_bands = std::max(n_bands, (int)(total / (1u << 30)) + 1);
    n_bands = std::max(1, std::min(n_bands, height));
    std::vector<PngBand> bands((size_t)n_bands);
    {
        std::vector<std::thread> th;
        for (int b = 0; b < n_bands; ++b) {
            const int y0 = (int)((long long)height * b / n_bands), y1 = (int)((long long)height * (b + 1) / n_bands);
            if (b + 1 < n_bands) th.emplace_back(deflate_band, rgb8, stride, y0, y1, false, &bands[(size_t)b]);
            else deflate_band(rgb8, stride, y0, y1, true, &bands[(size_t)b]);
        }
        for (auto& t : th) t.join();
    }
    std::vector<unsigned char> z;
    z.push_back(0x78); z.push_back(0x01);                             // zlib header: deflate, 32 KB window, fastest
    uLong adler = adler32(0L, Z_NULL, 0);
    for (const PngBand& b : bands) {
        if (!b.ok) { rt_set_error("rt_write_png_rgb8: deflate failed"); return RT_ERR_IO; }
        z.insert(z.end(), b.z.begin(), b.z.end());
        adler = adler32_combine(adler, b.adler, (z_off_t)b.raw_bytes);
    }
    for (int k = 3; k >= 0; --k) z.push_back((unsigned char)(adler >> (8 * k)));
    const size_t zlen = z.size();
    if (zlen > 0x7fffffffu) { rt_set_error("rt_write_png_rgb8: image too large for one IDAT chunk"); return RT_ERR_IO; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { rt_set_error(std::string("Failed to write PNG: ") + path); return RT_ERR_IO; }
    auto be32 = [](unsigned char* p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; };
    bool ok = true;
    auto chunk = [&](const char* type, const unsigned char* data, size_t n) {
        unsigned char hdr[8]; be32(hdr, (uint32_t)n); std::memcpy(hdr + 4, type, 4);
        uLong crc = crc32(0L, hdr + 4, 4);
        if (n) crc = crc32(crc, data, (uInt)n);
        unsigned char tail[4]; be32(tail, (uint32_t)crc);
        ok = ok && std::fwrite(hdr, 1, 8, f) == 8 && (n == 0 || std::fwrite(data, 1, n, f) == n) && std::fwrite(tail, 1, 4, f) == 4;
    };
    const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    ok = std::fwrite(sig, 1, 8, f) == 8;
    unsigned char ihdr[13]; be32(ihdr, (uint32_t)width); be32(ihdr + 4, (uint32_t)height);
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", z.data(), zlen);
    chunk("IEND", nullptr, 0);
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { rt_set_error(std::string("Failed to write PNG: ") + path); return RT_ERR_IO; }
    return RT_OK;
}

}  // extern "C"
