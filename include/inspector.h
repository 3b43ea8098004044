/* inspector.h -- headless stand-in for the reference's viewer (visualization/inspector.h:6-8):
 *
 *     typedef void (*Callback)(void* callbackData, unsigned char** image, int* out_width, int* out_height);
 *     int renderLoop(Callback callback, void* callbackData);
 *
 * The reference's renderLoop (visualization/inspector.cpp:71-282) opens a GLFW / OpenGL window, calls the callback
 * once per displayed frame and uploads the RGBA8 image it returns to a texture. A compute box has no window system
 * (unavailable on this box), so this loop calls the same callback for a fixed number of frames and writes the images
 * it returns as PNG files. Frame count and output come from RD::inspector::config() or the environment
 * (RD_INSPECTOR_FRAMES, RD_INSPECTOR_OUTPUT = printf pattern with one %d, RD_INSPECTOR_EVERY). The callback protocol
 * -- the image buffer belongs to the caller, the callback fills the three out-parameters -- is unchanged, so the
 * `render` function of samples/sample1.cpp:447-499 runs as it is. */
#ifndef RADIANCE_B200_INSPECTOR_H
#define RADIANCE_B200_INSPECTOR_H

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

typedef void (*Callback)(void* callbackData, unsigned char** image, int* out_width, int* out_height);

namespace RD
{
namespace inspector
{

struct Config { int frames = 1; int every = 1; std::string output; };   // output: printf pattern with one %d ("" = no files)
inline Config& config()
{
    static Config c = [] {
        Config k;
        if (const char* e = getenv("RD_INSPECTOR_FRAMES")) k.frames = atoi(e);
        if (const char* e = getenv("RD_INSPECTOR_EVERY")) k.every = atoi(e) > 0 ? atoi(e) : 1;
        if (const char* e = getenv("RD_INSPECTOR_OUTPUT")) k.output = e;
        return k;
    }();
    return c;
}

// PNG with stored (uncompressed) deflate blocks: enough for a frame dump, no dependency
inline bool writePng(const char* path, const unsigned char* rgba, int width, int height)
{
    static uint32_t table[256]; static bool ready = false;
    if (!ready) { for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; } ready = true; }
    auto crc = [&](const std::vector<unsigned char>& d, size_t from) { uint32_t c = 0xffffffffu; for (size_t i = from; i < d.size(); ++i) c = table[(c ^ d[i]) & 0xff] ^ (c >> 8); return c ^ 0xffffffffu; };
    auto be32 = [](std::vector<unsigned char>& d, uint32_t v) { d.push_back((unsigned char)(v >> 24)); d.push_back((unsigned char)(v >> 16)); d.push_back((unsigned char)(v >> 8)); d.push_back((unsigned char)v); };
    std::vector<unsigned char> raw;
    raw.reserve((size_t)height * ((size_t)width * 4 + 1));
    for (int y = 0; y < height; ++y) { raw.push_back(0); raw.insert(raw.end(), rgba + (size_t)y * width * 4, rgba + (size_t)(y + 1) * width * 4); }
    std::vector<unsigned char> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (size_t at = 0; at < raw.size() || at == 0; at += 65535) {
        const size_t len = raw.size() - at < 65535 ? raw.size() - at : 65535;
        z.push_back(at + len >= raw.size() ? 1 : 0);
        z.push_back((unsigned char)(len & 0xff)); z.push_back((unsigned char)(len >> 8));
        z.push_back((unsigned char)(~len & 0xff)); z.push_back((unsigned char)((~len >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + at, raw.begin() + at + len);
        if (at + len >= raw.size()) break;
    }
    for (unsigned char c : raw) { a = (a + c) % 65521u; b = (b + a) % 65521u; }
    be32(z, (b << 16) | a);
    std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    auto chunk = [&](const char* tag, const std::vector<unsigned char>& body) {
        be32(out, (uint32_t)body.size());
        const size_t from = out.size();
        out.insert(out.end(), tag, tag + 4);
        out.insert(out.end(), body.begin(), body.end());
        be32(out, crc(out, from));
    };
    std::vector<unsigned char> ihdr;
    be32(ihdr, (uint32_t)width); be32(ihdr, (uint32_t)height);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk("IHDR", ihdr); chunk("IDAT", z); chunk("IEND", {});
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
    fclose(f);
    return ok;
}

}  // namespace inspector
}  // namespace RD

inline int renderLoop(Callback callback, void* callbackData)
{
    const RD::inspector::Config& c = RD::inspector::config();
    for (int k = 0; k < c.frames; ++k) {
        unsigned char* image = nullptr; int width = 0, height = 0;
        callback(callbackData, &image, &width, &height);
        if (!c.output.empty() && (k % c.every == 0 || k == c.frames - 1)) {
            if (!image || width <= 0 || height <= 0) { printf("renderLoop: the callback returned no image\n"); return -1; }
            char path[1024];
            snprintf(path, sizeof(path), c.output.c_str(), k);
            if (!RD::inspector::writePng(path, image, width, height)) { printf("renderLoop: cannot write %s\n", path); return -1; }
        }
    }
    return 0;
}

#endif
