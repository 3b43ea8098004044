/* sceneBuilder.h -- `RD::Scene::Load` for the namespace-RD shim (include/radiance.h): the interface of the reference's
 * tools/sceneBuilder.h:8-56 (INCLUDE_SCENE_DESC, INCLUDE_SCENE_LAYOUT, struct RD::Scene) without assimp and stb.
 *
 * The reference reads assets with assimp (tools/sceneBuilder.cpp:27-36) and decodes embedded textures with stb_image
 * (:49-67). Neither is a dependency here: glTF 2.0 / GLB is parsed directly (a JSON document plus binary buffers) and
 * the conventions of assimp's glTF2 importer that the 9-descriptor bundle depends on are restated -- one mesh per
 * primitive, a default material appended to the list, TEXCOORD_0 with v flipped, embedded images as "*N", node transform
 * = matrix or T*R*S, several scene roots under an identity root, depth-first instance order (file:line references in
 * radiance-ray-tracing_b200/scene_loader.py, the Python twin of this header, which the tests compare it with).
 * Embedded PNG images are decoded by the small inflate below; other formats need `RD::Scene::decodeImage` to be set.
 * `meshInfoData` has one row per instance (see scene_loader.py for why). Errors abort like the rest of the shim. */
#ifndef RADIANCE_B200_SCENE_BUILDER_H
#define RADIANCE_B200_SCENE_BUILDER_H

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "radiance.h"

#define INCLUDE_SCENE_DESC(scene)   \
scene->meshInfoData,                \
scene->vertexData,                  \
scene->indexData,                   \
scene->uvData,                      \
scene->normalData,                  \
scene->materialData,                \
scene->textureData,                 \
scene->sampler,                     \
scene->topAccelStruct

#define INCLUDE_SCENE_LAYOUT        \
RD::BUFFER_TYPE,                    \
RD::BUFFER_TYPE,                    \
RD::BUFFER_TYPE,                    \
RD::BUFFER_TYPE,                    \
RD::BUFFER_TYPE,                    \
RD::BUFFER_TYPE,                    \
RD::TEX_ARRAY_TYPE,                 \
RD::IMAGE_SAMPLER_TYPE,             \
RD::ACCEL_STRUCT_TYPE

namespace RD
{
namespace gltf
{

[[noreturn]] inline void fail(const std::string& what)
{
    printf("RD::Scene::Load: %s\n", what.c_str());
    exit(-1);
}

// ---- a JSON value (just what glTF needs) ------------------------------------------------------------
struct Json
{
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    double number = 0;
    bool boolean = false;
    std::string text;
    std::vector<Json> items;
    std::vector<std::pair<std::string, Json>> members;

    const Json* find(const char* key) const
    {
        for (auto& m : members) if (m.first == key) return &m.second;
        return nullptr;
    }
    bool has(const char* key) const { return find(key) != nullptr; }
    const Json& at(const char* key) const
    {
        const Json* j = find(key);
        if (!j) fail(std::string("missing JSON member \"") + key + "\"");
        return *j;
    }
    const Json& at(int i) const { return at((size_t)i); }
    const Json& at(size_t i) const
    {
        if (kind != Array || i >= items.size()) fail("JSON array index out of range");
        return items[i];
    }
    size_t size() const { return kind == Array ? items.size() : members.size(); }
    double num(const char* key, double fallback) const { const Json* j = find(key); return j && j->kind == Number ? j->number : fallback; }
    long integer(const char* key, long fallback) const { return (long)num(key, (double)fallback); }
};

struct JsonParser
{
    const char* p; const char* end;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p; }
    bool eat(char c) { ws(); if (p < end && *p == c) { ++p; return true; } return false; }
    void expect(char c) { if (!eat(c)) fail(std::string("malformed JSON: expected '") + c + "'"); }
    std::string string()
    {
        expect('"');
        std::string s;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                ++p;
                switch (*p) {
                case 'n': s += '\n'; break; case 't': s += '\t'; break; case 'r': s += '\r'; break;
                case 'b': s += '\b'; break; case 'f': s += '\f'; break;
                case 'u': {
                    if (p + 4 >= end) fail("malformed JSON: \\u escape");
                    unsigned cp = (unsigned)strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                    p += 4;
                    if (cp < 0x80) s += (char)cp;
                    else if (cp < 0x800) { s += (char)(0xC0 | (cp >> 6)); s += (char)(0x80 | (cp & 0x3F)); }
                    else { s += (char)(0xE0 | (cp >> 12)); s += (char)(0x80 | ((cp >> 6) & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: s += *p;
                }
                ++p;
            } else s += *p++;
        }
        if (p >= end) fail("malformed JSON: unterminated string");
        ++p;
        return s;
    }
    Json value()
    {
        ws();
        if (p >= end) fail("malformed JSON: unexpected end");
        Json j;
        if (*p == '{') {
            ++p; j.kind = Json::Object;
            if (eat('}')) return j;
            do { std::string k = string(); expect(':'); j.members.emplace_back(k, value()); } while (eat(','));
            expect('}');
        } else if (*p == '[') {
            ++p; j.kind = Json::Array;
            if (eat(']')) return j;
            do { j.items.push_back(value()); } while (eat(','));
            expect(']');
        } else if (*p == '"') { j.kind = Json::String; j.text = string(); }
        else if (!strncmp(p, "true", 4) && end - p >= 4) { j.kind = Json::Bool; j.boolean = true; p += 4; }
        else if (!strncmp(p, "false", 5) && end - p >= 5) { j.kind = Json::Bool; p += 5; }
        else if (!strncmp(p, "null", 4) && end - p >= 4) { p += 4; }
        else {
            char* stop = nullptr;
            j.kind = Json::Number; j.number = strtod(p, &stop);
            if (stop == p) fail("malformed JSON: unexpected character");
            p = stop;
        }
        return j;
    }
};

inline std::vector<unsigned char> base64(const std::string& s, size_t from)
{
    std::vector<unsigned char> out;
    unsigned acc = 0; int bits = 0;
    for (size_t i = from; i < s.size(); ++i) {
        const char c = s[i];
        int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A'; else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
        else if (c >= '0' && c <= '9') v = c - '0' + 52; else if (c == '+' || c == '-') v = 62; else if (c == '/' || c == '_') v = 63;
        else continue;
        acc = (acc << 6) | (unsigned)v; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((unsigned char)((acc >> bits) & 0xff)); }
    }
    return out;
}

inline std::vector<unsigned char> readFile(const std::string& path)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) fail("cannot open " + path);
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<unsigned char> d((size_t)(n > 0 ? n : 0));
    if (n > 0 && fread(d.data(), 1, (size_t)n, f) != (size_t)n) { fclose(f); fail("short read on " + path); }
    fclose(f);
    return d;
}

struct Document
{
    Json doc;
    std::vector<std::vector<unsigned char>> buffers;
};

inline uint32_t u32le(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

inline Document readContainer(const std::string& path)
{
    Document D;
    std::vector<unsigned char> raw = readFile(path), binary;
    bool haveBinary = false;
    if (raw.size() >= 12 && !memcmp(raw.data(), "glTF", 4)) {
        if (u32le(&raw[4]) != 2) fail(path + ": GLB container version " + std::to_string(u32le(&raw[4])) + ", only 2 is supported");
        const size_t length = u32le(&raw[8]);
        if (length > raw.size()) fail(path + ": truncated GLB");
        size_t at = 12; bool haveJson = false;
        while (at + 8 <= length) {
            const size_t clen = u32le(&raw[at]); const uint32_t ctype = u32le(&raw[at + 4]);
            if (at + 8 + clen > raw.size()) fail(path + ": truncated GLB chunk");
            if (ctype == 0x4E4F534Au) { JsonParser P{(const char*)&raw[at + 8], (const char*)&raw[at + 8] + clen}; D.doc = P.value(); haveJson = true; }
            else if (ctype == 0x004E4942u && !haveBinary) { binary.assign(raw.begin() + at + 8, raw.begin() + at + 8 + clen); haveBinary = true; }
            at += 8 + clen + ((4 - clen % 4) % 4);
        }
        if (!haveJson) fail(path + ": GLB without a JSON chunk");
    } else {
        JsonParser P{(const char*)raw.data(), (const char*)raw.data() + raw.size()};
        D.doc = P.value();
    }
    const Json* bufs = D.doc.find("buffers");
    for (size_t k = 0; bufs && k < bufs->size(); ++k) {
        const Json& b = bufs->at(k);
        std::vector<unsigned char> data;
        const Json* uri = b.find("uri");
        if (!uri) { if (!haveBinary) fail(path + ": buffer without uri and no BIN chunk"); data = binary; }
        else if (uri->text.compare(0, 5, "data:") == 0) data = base64(uri->text, uri->text.find(',') + 1);
        else {
            const size_t slash = path.find_last_of("/\\");
            data = readFile((slash == std::string::npos ? std::string() : path.substr(0, slash + 1)) + uri->text);
        }
        if (data.size() < (size_t)b.integer("byteLength", 0)) fail(path + ": buffer " + std::to_string(k) + " shorter than its byteLength");
        D.buffers.push_back(std::move(data));
    }
    return D;
}

// ---- accessors -------------------------------------------------------------------------------------
// element (count x width) as double, honouring byteOffset / byteStride / normalized
struct Accessor { size_t count = 0; int width = 0; std::vector<double> v; bool isFloat = false; };

inline Accessor readAccessor(const Document& D, long index)
{
    const Json& a = D.doc.at("accessors").at((size_t)index);
    if (a.has("sparse")) fail("accessor " + std::to_string(index) + ": sparse accessors are not supported");
    const long ct = a.integer("componentType", 0);
    const std::string& type = a.at("type").text;
    Accessor A;
    A.count = (size_t)a.integer("count", 0);
    A.width = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : type == "MAT4" ? 16 : type == "MAT3" ? 9 : 4;
    const size_t comp = (ct == 5120 || ct == 5121) ? 1 : (ct == 5122 || ct == 5123) ? 2 : 4;
    A.isFloat = ct == 5126;
    A.v.assign(A.count * A.width, 0.0);
    if (!a.has("bufferView") || !A.count) return A;
    const Json& view = D.doc.at("bufferViews").at((size_t)a.integer("bufferView", 0));
    const size_t start = (size_t)view.integer("byteOffset", 0) + (size_t)a.integer("byteOffset", 0);
    const size_t elem = comp * A.width;
    const size_t stride = view.integer("byteStride", 0) ? (size_t)view.integer("byteStride", 0) : elem;
    const std::vector<unsigned char>& buf = D.buffers.at((size_t)view.integer("buffer", 0));
    if (start + stride * (A.count - 1) + elem > buf.size()) fail("accessor " + std::to_string(index) + ": reads past the end of its buffer");
    const bool normalized = a.find("normalized") && a.at("normalized").boolean;
    for (size_t i = 0; i < A.count; ++i)
        for (int c = 0; c < A.width; ++c) {
            const unsigned char* p = &buf[start + i * stride + c * comp];
            double x;
            switch (ct) {
            case 5120: { int8_t t; memcpy(&t, p, 1); x = t; if (normalized) x = std::max(x / 127.0, -1.0); break; }
            case 5121: { x = p[0]; if (normalized) x /= 255.0; break; }
            case 5122: { int16_t t; memcpy(&t, p, 2); x = t; if (normalized) x = std::max(x / 32767.0, -1.0); break; }
            case 5123: { uint16_t t; memcpy(&t, p, 2); x = t; if (normalized) x /= 65535.0; break; }
            case 5125: { uint32_t t; memcpy(&t, p, 4); x = t; break; }
            case 5126: { float t; memcpy(&t, p, 4); x = t; break; }
            default: fail("accessor " + std::to_string(index) + ": component type " + std::to_string(ct));
            }
            A.v[i * A.width + c] = x;
        }
    return A;
}

// ---- PNG (8-bit, non-interlaced) through a small inflate ---------------------------------------------
struct BitReader
{
    const unsigned char* p; size_t n, at = 0; uint32_t acc = 0; int have = 0;
    uint32_t bits(int k)
    {
        while (have < k) { if (at >= n) fail("PNG: deflate stream ends early"); acc |= (uint32_t)p[at++] << have; have += 8; }
        const uint32_t v = acc & ((k == 32) ? 0xffffffffu : ((1u << k) - 1u));
        acc >>= k; have -= k;
        return v;
    }
};

struct Huffman
{
    uint16_t count[16] = {0}; uint16_t symbol[320] = {0};
    void build(const uint8_t* lengths, int n)
    {
        memset(count, 0, sizeof(count));
        for (int i = 0; i < n; ++i) count[lengths[i]]++;
        count[0] = 0;
        uint16_t offs[16]; offs[1] = 0;
        for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
        for (int i = 0; i < n; ++i) if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
    }
    int decode(BitReader& br) const
    {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= 15; ++l) {
            code |= (int)br.bits(1);
            const int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        fail("PNG: bad Huffman code");
    }
};

inline std::vector<unsigned char> inflateZlib(const unsigned char* data, size_t n)
{
    if (n < 6) fail("PNG: zlib stream too short");
    BitReader br{data + 2, n - 2};
    std::vector<unsigned char> out;
    static const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const uint16_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    static const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const uint16_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    for (;;) {
        const uint32_t last = br.bits(1), type = br.bits(2);
        if (type == 0) {
            br.acc = 0; br.have = 0;
            if (br.at + 4 > br.n) fail("PNG: stored block header");
            const size_t len = br.p[br.at] | (br.p[br.at + 1] << 8);
            br.at += 4;
            if (br.at + len > br.n) fail("PNG: stored block overruns");
            out.insert(out.end(), br.p + br.at, br.p + br.at + len);
            br.at += len;
        } else if (type == 1 || type == 2) {
            Huffman lit, dist;
            uint8_t lengths[320];
            if (type == 1) {
                int i = 0;
                for (; i < 144; ++i) lengths[i] = 8;
                for (; i < 256; ++i) lengths[i] = 9;
                for (; i < 280; ++i) lengths[i] = 7;
                for (; i < 288; ++i) lengths[i] = 8;
                lit.build(lengths, 288);
                for (i = 0; i < 30; ++i) lengths[i] = 5;
                dist.build(lengths, 30);
            } else {
                const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
                static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                uint8_t cl[19] = {0};
                for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)br.bits(3);
                Huffman lencode; lencode.build(cl, 19);
                int i = 0;
                while (i < nlen + ndist) {
                    int sym = lencode.decode(br);
                    if (sym < 16) lengths[i++] = (uint8_t)sym;
                    else {
                        int rep, val = 0;
                        if (sym == 16) { if (!i) fail("PNG: bad code lengths"); val = lengths[i - 1]; rep = 3 + (int)br.bits(2); }
                        else if (sym == 17) rep = 3 + (int)br.bits(3);
                        else rep = 11 + (int)br.bits(7);
                        if (i + rep > nlen + ndist) fail("PNG: bad code lengths");
                        while (rep--) lengths[i++] = (uint8_t)val;
                    }
                }
                lit.build(lengths, nlen);
                dist.build(lengths + nlen, ndist);
            }
            for (;;) {
                int sym = lit.decode(br);
                if (sym < 256) out.push_back((unsigned char)sym);
                else if (sym == 256) break;
                else {
                    sym -= 257;
                    if (sym >= 29) fail("PNG: bad length symbol");
                    const size_t len = lbase[sym] + br.bits(lext[sym]);
                    const int ds = dist.decode(br);
                    if (ds >= 30) fail("PNG: bad distance symbol");
                    const size_t d = dbase[ds] + br.bits(dext[ds]);
                    if (d > out.size()) fail("PNG: distance before the start of the stream");
                    const size_t from = out.size() - d;
                    for (size_t k = 0; k < len; ++k) out.push_back(out[from + k]);
                }
            }
        } else fail("PNG: reserved deflate block type");
        if (last) break;
    }
    return out;
}

// PNG -> RGBA8 (w, h). Supports colour types 0 / 2 / 3 / 4 / 6 at 8 bits per channel, no interlace.
inline bool decodePng(const unsigned char* data, size_t n, std::vector<unsigned char>& rgba, int& w, int& h)
{
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 8 || memcmp(data, sig, 8)) return false;
    auto be32 = [](const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3]; };
    std::vector<unsigned char> idat, palette, trns;
    int depth = 0, ctype = 0, interlace = 0;
    size_t at = 8;
    while (at + 12 <= n) {
        const size_t len = be32(data + at);
        const unsigned char* tag = data + at + 4; const unsigned char* body = data + at + 8;
        if (at + 12 + len > n) fail("PNG: truncated chunk");
        if (!memcmp(tag, "IHDR", 4)) { w = (int)be32(body); h = (int)be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12]; }
        else if (!memcmp(tag, "PLTE", 4)) palette.assign(body, body + len);
        else if (!memcmp(tag, "tRNS", 4)) trns.assign(body, body + len);
        else if (!memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!memcmp(tag, "IEND", 4)) break;
        at += 12 + len;
    }
    if (depth != 8 || interlace) fail("PNG: only 8-bit non-interlaced images are supported by the built-in decoder (set RD::Scene::decodeImage)");
    const int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
    std::vector<unsigned char> raw = inflateZlib(idat.data(), idat.size());
    const size_t stride = (size_t)w * ch;
    if (raw.size() < (stride + 1) * (size_t)h) fail("PNG: image data too short");
    std::vector<unsigned char> px(stride * h);
    for (int y = 0; y < h; ++y) {
        const unsigned char* src = &raw[(stride + 1) * y];
        unsigned char* dst = &px[stride * y];
        const unsigned char* up = y ? dst - stride : nullptr;
        const int filter = src[0];
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= (size_t)ch ? dst[x - ch] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
            int v = src[1 + x];
            switch (filter) {
            case 0: break;
            case 1: v += a; break;
            case 2: v += b; break;
            case 3: v += (a + b) / 2; break;
            case 4: { const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c); v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
            default: fail("PNG: bad filter type");
            }
            dst[x] = (unsigned char)v;
        }
    }
    rgba.resize((size_t)w * h * 4);
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        unsigned char r, g, b, a = 255;
        const unsigned char* s = &px[i * ch];
        if (ctype == 0) r = g = b = s[0];
        else if (ctype == 2) { r = s[0]; g = s[1]; b = s[2]; }
        else if (ctype == 3) { const size_t k = s[0]; if (3 * k + 2 >= palette.size()) fail("PNG: palette index"); r = palette[3 * k]; g = palette[3 * k + 1]; b = palette[3 * k + 2]; if (k < trns.size()) a = trns[k]; }
        else if (ctype == 4) { r = g = b = s[0]; a = s[1]; }
        else { r = s[0]; g = s[1]; b = s[2]; a = s[3]; }
        rgba[4 * i] = r; rgba[4 * i + 1] = g; rgba[4 * i + 2] = b; rgba[4 * i + 3] = a;
    }
    return true;
}

// Resize of RGBA8 as stbir_resize_uint8_linear(..., STBIR_RGBA) does it for sceneBuilder.cpp:60-62 -- stb_image_resize2's default
// filters: Catmull-Rom when enlarging, Mitchell-Netravali (B = C = 1/3) widened by 1 / scale when shrinking; edge clamp, texel
// centres at i + 0.5, weights normalised, two separable passes. Checked against the reference's own stb code through the Python
// twin of this function (scene_loader.decode_texture): identical for 16 -> 4096, within one level on 1e-5 of the texels for
// 8192 -> 4096. Alpha is taken as opaque for the filter.
inline float resizeKernel(float x, bool upsample)
{
    x = std::fabs(x);
    if (upsample) return x < 1.0f ? 1.0f - x * x * (2.5f - 1.5f * x) : (x < 2.0f ? 2.0f - x * (4.0f + x * (0.5f * x - 2.5f)) : 0.0f);
    return x < 1.0f ? (16.0f + x * x * (21.0f * x - 36.0f)) / 18.0f : (x < 2.0f ? (32.0f + x * (-60.0f + x * (36.0f - 7.0f * x))) / 18.0f : 0.0f);
}

struct ResizeTaps { std::vector<int> first; std::vector<float> weight; int taps; };   // per output texel: first input texel, `taps` normalised weights
inline ResizeTaps resizeTaps(int nIn, int nOut)
{
    ResizeTaps T;
    const double scale = (double)nOut / (double)nIn;
    const bool upsample = scale >= 1.0;
    const double support = upsample ? 2.0 : 2.0 / scale, kscale = upsample ? 1.0 : scale;
    T.taps = (int)std::ceil(2.0 * support) + 1;
    T.first.resize((size_t)nOut); T.weight.resize((size_t)nOut * T.taps);
    for (int o = 0; o < nOut; ++o) {
        const double centre = ((double)o + 0.5) / scale;
        const int lo = (int)std::floor(centre - support + 0.5);
        T.first[(size_t)o] = lo;
        double sum = 0.0;
        for (int k = 0; k < T.taps; ++k) { const float w = resizeKernel((float)((centre - ((double)(lo + k) + 0.5)) * kscale), upsample); T.weight[(size_t)o * T.taps + k] = w; sum += w; }
        for (int k = 0; k < T.taps; ++k) T.weight[(size_t)o * T.taps + k] = (float)(T.weight[(size_t)o * T.taps + k] / sum);
    }
    return T;
}

inline std::vector<unsigned char> resizeRgba(const std::vector<unsigned char>& src, int w, int h, int dim)
{
    if (w == dim && h == dim) return src;
    // horizontal pass into floats (h rows of dim texels), then the vertical pass
    std::vector<float> wide((size_t)h * dim * 4);
    if (w == dim) for (size_t i = 0; i < wide.size(); ++i) wide[i] = (float)src[i];
    else {
        const ResizeTaps X = resizeTaps(w, dim);
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < dim; ++x) {
                float acc[4] = {0, 0, 0, 0};
                for (int k = 0; k < X.taps; ++k) {
                    int i = X.first[(size_t)x] + k; i = i < 0 ? 0 : (i > w - 1 ? w - 1 : i);
                    const float wgt = X.weight[(size_t)x * X.taps + k];
                    const unsigned char* p = &src[((size_t)y * w + i) * 4];
                    acc[0] += wgt * p[0]; acc[1] += wgt * p[1]; acc[2] += wgt * p[2]; acc[3] += wgt * p[3];
                }
                float* q = &wide[((size_t)y * dim + x) * 4];
                q[0] = acc[0]; q[1] = acc[1]; q[2] = acc[2]; q[3] = acc[3];
            }
    }
    std::vector<unsigned char> out((size_t)dim * dim * 4);
    auto store = [](float v) { v = std::floor(v + 0.5f); return (unsigned char)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v)); };
    if (h == dim) { for (size_t i = 0; i < out.size(); ++i) out[i] = store(wide[i]); return out; }
    const ResizeTaps Y = resizeTaps(h, dim);
    for (int y = 0; y < dim; ++y)
        for (int x = 0; x < dim; ++x) {
            float acc[4] = {0, 0, 0, 0};
            for (int k = 0; k < Y.taps; ++k) {
                int i = Y.first[(size_t)y] + k; i = i < 0 ? 0 : (i > h - 1 ? h - 1 : i);
                const float wgt = Y.weight[(size_t)y * Y.taps + k];
                const float* p = &wide[((size_t)i * dim + x) * 4];
                acc[0] += wgt * p[0]; acc[1] += wgt * p[1]; acc[2] += wgt * p[2]; acc[3] += wgt * p[3];
            }
            unsigned char* q = &out[((size_t)y * dim + x) * 4];
            q[0] = store(acc[0]); q[1] = store(acc[1]); q[2] = store(acc[2]); q[3] = store(acc[3]);
        }
    return out;
}

// ---- the asset as host arrays (the layout RD::Scene uploads) ---------------------------------------
struct Parsed
{
    struct MeshData { std::vector<Vec3> vertices; std::vector<Triangle> triangles; };
    struct InstanceData { unsigned mesh; Mat4x4 transform; unsigned SBTOffset, customInstanceID; };
    std::vector<MeshData> meshes;
    std::vector<MeshInfo> meshInfoPerMesh, meshInfo;      // per mesh (reference) / per instance (uploaded)
    std::vector<Vec3> vertices, uvs, normals;
    std::vector<Triangle> triangles;
    std::vector<Material> materials;
    std::vector<InstanceData> instances;
    std::vector<std::vector<unsigned char>> images;       // encoded, in "*N" order
};

inline Mat4x4 multiply(const Mat4x4& a, const Mat4x4& b)
{
    Mat4x4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            float s = 0;
            for (int k = 0; k < 4; ++k) s += a[i][k] * b[k][j];
            r[i][j] = s;
        }
    return r;
}

inline Mat4x4 nodeTransform(const Json& node)
{
    Mat4x4 m;
    if (const Json* mat = node.find("matrix")) {          // glTF stores column-major
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m[r][c] = (float)mat->at((size_t)(c * 4 + r)).number;
        return m;
    }
    if (const Json* t = node.find("translation")) {
        Mat4x4 tm; tm[0][3] = (float)t->at(0).number; tm[1][3] = (float)t->at(1).number; tm[2][3] = (float)t->at(2).number;
        m = multiply(m, tm);
    }
    if (const Json* q = node.find("rotation")) {
        const double x = q->at(0).number, y = q->at(1).number, z = q->at(2).number, w = q->at(3).number;
        Mat4x4 r;
        r[0][0] = (float)(1 - 2 * (y * y + z * z)); r[0][1] = (float)(2 * (x * y - z * w)); r[0][2] = (float)(2 * (x * z + y * w));
        r[1][0] = (float)(2 * (x * y + z * w)); r[1][1] = (float)(1 - 2 * (x * x + z * z)); r[1][2] = (float)(2 * (y * z - x * w));
        r[2][0] = (float)(2 * (x * z - y * w)); r[2][1] = (float)(2 * (y * z + x * w)); r[2][2] = (float)(1 - 2 * (x * x + y * y));
        m = multiply(m, r);
    }
    if (const Json* s = node.find("scale")) {
        Mat4x4 sm; sm[0][0] = (float)s->at(0).number; sm[1][1] = (float)s->at(1).number; sm[2][2] = (float)s->at(2).number;
        m = multiply(m, sm);
    }
    return m;
}

// aiProcess_GenSmoothNormals as assimp runs it on the importer's INDEXED mesh (GenVertexNormalsProcess.cpp:104-235, default
// smoothing angle 175 degrees): every face writes its unit normal ((v1 - v0) x (v2 - v0), NormalizeSafe) to its three vertices
// -- for a vertex shared by several faces the LAST face in index order wins -- and vertices at the same position then share the
// normalised sum of those per-vertex normals (assimp groups positions within 1e-4 of the mesh's extent; exact equality here).
inline std::vector<Vec3> importerNormals(const std::vector<Vec3>& v, const std::vector<Triangle>& t)
{
    std::vector<Vec3> perVertex(v.size(), Vec3(0.0f, 0.0f, 0.0f));
    for (const Triangle& f : t) {
        const Vec3 &a = v[f.idx0], &b = v[f.idx1], &c = v[f.idx2];
        const float e1[3] = {b.x - a.x, b.y - a.y, b.z - a.z}, e2[3] = {c.x - a.x, c.y - a.y, c.z - a.z};
        float n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        const float len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        if (len > 0.0f) { n[0] /= len; n[1] /= len; n[2] /= len; }
        for (unsigned idx : {f.idx0, f.idx1, f.idx2}) perVertex[idx] = Vec3(n[0], n[1], n[2]);
    }
    struct Key { float x, y, z; bool operator<(const Key& o) const { return x != o.x ? x < o.x : (y != o.y ? y < o.y : z < o.z); } };
    std::map<Key, size_t> groupOf;
    std::vector<size_t> group(v.size());
    for (size_t i = 0; i < v.size(); ++i) group[i] = groupOf.emplace(Key{v[i].x, v[i].y, v[i].z}, groupOf.size()).first->second;
    std::vector<float> acc(groupOf.size() * 3, 0.0f);
    for (size_t i = 0; i < v.size(); ++i) { acc[group[i] * 3] += perVertex[i].x; acc[group[i] * 3 + 1] += perVertex[i].y; acc[group[i] * 3 + 2] += perVertex[i].z; }
    std::vector<Vec3> out(v.size());
    for (size_t i = 0; i < v.size(); ++i) {
        const float* n = &acc[group[i] * 3];
        const float len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        out[i] = len > 0.0f ? Vec3(n[0] / len, n[1] / len, n[2] / len) : Vec3(n[0], n[1], n[2]);
    }
    return out;
}

// The vertex numbering assimp hands to sceneBuilder.cpp: the glTF2 importer keeps only the vertices the index list uses and
// numbers them in the order the index list first mentions them (glTF2Importer.cpp ImportMeshes, vertexRemappingTable); missing
// normals are generated on that mesh; then aiProcess_JoinIdenticalVertices merges vertices whose position, normal and uv are
// identical, keeping the first (JoinVerticesProcess.cpp:258-381; its comparison has a 1e-5 tolerance, exact equality is used
// here). Triangles keep their order. `normals` empty = generate them.
inline void importerVertexOrder(std::vector<Vec3>& v, std::vector<Triangle>& t, std::vector<Vec3>& normals, std::vector<Vec3>& uv)
{
    const unsigned none = 0xffffffffu;
    std::vector<unsigned> remap(v.size(), none), order;
    for (Triangle& f : t)
        for (unsigned* idx : {&f.idx0, &f.idx1, &f.idx2}) {
            if (remap[*idx] == none) { remap[*idx] = (unsigned)order.size(); order.push_back(*idx); }
            *idx = remap[*idx];
        }
    std::vector<Vec3> v2(order.size()), uv2(order.size()), n2;
    for (size_t i = 0; i < order.size(); ++i) { v2[i] = v[order[i]]; uv2[i] = uv[order[i]]; }
    if (!normals.empty()) { n2.resize(order.size()); for (size_t i = 0; i < order.size(); ++i) n2[i] = normals[order[i]]; }
    else n2 = importerNormals(v2, t);
    struct Key { float a[9]; bool operator<(const Key& o) const { return memcmp(a, o.a, sizeof(a)) < 0; } };
    std::map<Key, unsigned> firstOf;
    std::vector<unsigned> joined(order.size());
    std::vector<Vec3> v3, n3, uv3;
    for (size_t i = 0; i < order.size(); ++i) {
        Key k = {{v2[i].x, v2[i].y, v2[i].z, n2[i].x, n2[i].y, n2[i].z, uv2[i].x, uv2[i].y, uv2[i].z}};
        for (float& x : k.a) if (x == 0.0f) x = 0.0f;   // -0 and +0 are one value
        auto it = firstOf.find(k);
        if (it == firstOf.end()) { it = firstOf.emplace(k, (unsigned)v3.size()).first; v3.push_back(v2[i]); n3.push_back(n2[i]); uv3.push_back(uv2[i]); }
        joined[i] = it->second;
    }
    for (Triangle& f : t) { f.idx0 = joined[f.idx0]; f.idx1 = joined[f.idx1]; f.idx2 = joined[f.idx2]; }
    v.swap(v3); normals.swap(n3); uv.swap(uv3);
}

inline int textureImage(const Json& doc, const std::vector<int>& embedded, const Json* info)
{
    if (!info) return -1;
    const Json& tex = doc.at("textures").at((size_t)info->integer("index", 0));
    const long image = tex.integer("source", -1);
    if (image < 0 || (size_t)image >= embedded.size() || embedded[(size_t)image] < 0)
        fail("texture " + std::to_string(info->integer("index", 0)) + ": only embedded images are supported");
    return embedded[(size_t)image];
}

inline Parsed parse(const std::string& path)
{
    const Document D = readContainer(path);
    const Json& doc = D.doc;
    Parsed S;
    // assimp's glTF2 importer loads only what the default scene references, in the order it first meets it (LazyDict::Retrieve):
    // Node::Read (glTF2Asset.inl:1692-1760) retrieves a node's children -- depth first -- BEFORE the node's own mesh, Mesh::Read
    // retrieves each primitive's material, Material::Read (:1283-1300) the baseColor, metallicRoughness, normal, occlusion and
    // emissive textures and through them the images. aiMesh / aiMaterial / "*N" numbering follows that order, not the file's
    // (pinned against the reference's own loader on the reference's assimp by tests/test_scene_loader.py).
    std::vector<size_t> meshOrder, materialOrder, imageOrder;
    {
        const Json* nodesJ = doc.find("nodes");
        const Json* meshesJ = doc.find("meshes");
        const Json* matsJ = doc.find("materials");
        const Json* texJ = doc.find("textures");
        std::vector<char> seen(nodesJ ? nodesJ->size() : 0, 0);
        auto listed = [](const std::vector<size_t>& v, size_t x) { for (size_t y : v) if (y == x) return true; return false; };
        struct Loader
        {
            const Json *nodesJ, *meshesJ, *matsJ, *texJ; std::vector<char>& seen;
            std::vector<size_t>&meshOrder, &materialOrder, &imageOrder;
            bool (*listed)(const std::vector<size_t>&, size_t);
            void texture(const Json* info)
            {
                if (!info || !texJ) return;
                const long source = texJ->at((size_t)info->integer("index", 0)).integer("source", -1);
                if (source >= 0 && !listed(imageOrder, (size_t)source)) imageOrder.push_back((size_t)source);
            }
            void material(size_t k)
            {
                if (listed(materialOrder, k)) return;
                materialOrder.push_back(k);
                const Json& m = matsJ->at(k);
                const Json* pbr = m.find("pbrMetallicRoughness");
                texture(pbr ? pbr->find("baseColorTexture") : nullptr);
                texture(pbr ? pbr->find("metallicRoughnessTexture") : nullptr);
                texture(m.find("normalTexture")); texture(m.find("occlusionTexture")); texture(m.find("emissiveTexture"));
            }
            void node(size_t i)
            {
                if (i >= seen.size()) fail("node " + std::to_string(i) + ": index out of range");
                if (seen[i]) return;
                seen[i] = 1;
                const Json& n = nodesJ->at(i);
                if (const Json* ch = n.find("children")) for (size_t k = 0; k < ch->size(); ++k) node((size_t)ch->at(k).number);
                if (n.has("mesh")) {
                    const size_t mi = (size_t)n.integer("mesh", 0);
                    if (!meshesJ || mi >= meshesJ->size()) fail("node " + std::to_string(i) + ": mesh index out of range");
                    if (!listed(meshOrder, mi)) {
                        meshOrder.push_back(mi);
                        const Json& prims = meshesJ->at(mi).at("primitives");
                        for (size_t pi = 0; pi < prims.size(); ++pi) if (prims.at(pi).has("material")) material((size_t)prims.at(pi).integer("material", 0));
                    }
                }
            }
        } loader{nodesJ, meshesJ, matsJ, texJ, seen, meshOrder, materialOrder, imageOrder, +listed};
        const Json* scenesJ = doc.find("scenes");
        if (scenesJ && scenesJ->size() && nodesJ)
            if (const Json* roots = scenesJ->at((size_t)doc.integer("scene", 0)).find("nodes"))
                for (size_t k = 0; k < roots->size(); ++k) loader.node((size_t)roots->at(k).number);
    }
    // embedded images, in the order the importer meets them -> "*N"
    std::vector<int> embedded(doc.find("images") ? doc.find("images")->size() : 0, -1);
    for (size_t k : imageOrder) {
        const Json& img = doc.at("images").at(k);
        if (img.has("bufferView")) {
            const Json& view = doc.at("bufferViews").at((size_t)img.integer("bufferView", 0));
            const std::vector<unsigned char>& buf = D.buffers.at((size_t)view.integer("buffer", 0));
            const size_t at = (size_t)view.integer("byteOffset", 0), len = (size_t)view.integer("byteLength", 0);
            if (at + len > buf.size()) fail("image " + std::to_string(k) + ": buffer view outside its buffer");
            S.images.emplace_back(buf.begin() + at, buf.begin() + at + len);
        } else if (img.has("uri") && img.at("uri").text.compare(0, 5, "data:") == 0) {
            S.images.push_back(base64(img.at("uri").text, img.at("uri").text.find(',') + 1));
        } else continue;
        embedded[k] = (int)S.images.size() - 1;
    }
    // the materials the scene references + the importer's default material; the transmission / ior extensions count only when
    // the asset declares them in extensionsUsed (glTF2Asset.inl:1360, :1384)
    bool useTransmission = false, useIor = false;
    if (const Json* used = doc.find("extensionsUsed"))
        for (size_t k = 0; k < used->size(); ++k) {
            if (used->at(k).text == "KHR_materials_transmission") useTransmission = true;
            if (used->at(k).text == "KHR_materials_ior") useIor = true;
        }
    const Json* mats = doc.find("materials");
    const size_t nm = materialOrder.size();
    static const Json empty;
    for (size_t k = 0; k <= nm; ++k) {
        const Json& m = k < nm ? mats->at(materialOrder[k]) : empty;
        const Json* pbr = m.find("pbrMetallicRoughness");
        const Json* ext = m.find("extensions");
        Material M;
        for (int c = 0; c < 4; ++c) M.albedo[c] = 1.0f;
        if (pbr && pbr->has("baseColorFactor")) for (int c = 0; c < 4; ++c) M.albedo[c] = (float)pbr->at("baseColorFactor").at((size_t)c).number;
        M.albedoTexIdx = textureImage(doc, embedded, pbr ? pbr->find("baseColorTexture") : nullptr);
        M.metallic = (float)(pbr ? pbr->num("metallicFactor", 1.0) : 1.0);
        M.roughness = (float)(pbr ? pbr->num("roughnessFactor", 1.0) : 1.0);
        M.metallicTexIdx = M.roughnessTexIdx = textureImage(doc, embedded, pbr ? pbr->find("metallicRoughnessTexture") : nullptr);
        const Json* tr = (ext && useTransmission) ? ext->find("KHR_materials_transmission") : nullptr;
        const Json* ior = (ext && useIor) ? ext->find("KHR_materials_ior") : nullptr;
        M.transmission = (float)(tr ? tr->num("transmissionFactor", 0.0) : 0.0);
        M.ior = (float)(ior ? ior->num("ior", 1.45) : 1.45);
        M.normalTexIdx = textureImage(doc, embedded, m.find("normalTexture"));
        S.materials.push_back(M);
    }
    const int defaultMaterial = (int)nm;
    // one mesh per primitive
    const Json* meshes = doc.find("meshes");
    std::vector<size_t> meshFirst(meshes ? meshes->size() : 0, 0), meshEnd(meshes ? meshes->size() : 0, 0);   // per glTF mesh: its run of primitives
    for (size_t mi : meshOrder) {
        meshFirst[mi] = S.meshes.size();
        const Json& prims = meshes->at(mi).at("primitives");
        for (size_t pi = 0; pi < prims.size(); ++pi) {
            const Json& prim = prims.at(pi);
            const std::string where = "mesh " + std::to_string(mi) + " primitive " + std::to_string(pi);
            if (prim.integer("mode", 4) != 4) fail(where + ": mode " + std::to_string(prim.integer("mode", 4)) + "; only triangle lists (4) reach the tracer");
            const Json& attr = prim.at("attributes");
            if (!attr.has("POSITION")) fail(where + ": no POSITION attribute");
            const Accessor P = readAccessor(D, attr.integer("POSITION", 0));
            if (P.width != 3) fail(where + ": POSITION is not VEC3");
            Parsed::MeshData M;
            for (size_t i = 0; i < P.count; ++i) M.vertices.emplace_back((float)P.v[3 * i], (float)P.v[3 * i + 1], (float)P.v[3 * i + 2]);
            std::vector<unsigned> idx;
            if (prim.has("indices")) { const Accessor I = readAccessor(D, prim.integer("indices", 0)); for (double x : I.v) idx.push_back((unsigned)x); }
            else for (size_t i = 0; i < P.count; ++i) idx.push_back((unsigned)i);
            if (idx.size() % 3) fail(where + ": " + std::to_string(idx.size()) + " indices is not a multiple of 3");
            for (size_t i = 0; i < idx.size(); i += 3) {
                if (idx[i] >= P.count || idx[i + 1] >= P.count || idx[i + 2] >= P.count) fail(where + ": index out of range");
                M.triangles.push_back(Triangle{idx[i], idx[i + 1], idx[i + 2]});
            }
            std::vector<Vec3> normals;   // empty = generated by importerVertexOrder
            if (attr.has("NORMAL")) { const Accessor N = readAccessor(D, attr.integer("NORMAL", 0)); for (size_t i = 0; i < N.count; ++i) normals.emplace_back((float)N.v[3 * i], (float)N.v[3 * i + 1], (float)N.v[3 * i + 2]); if (normals.size() != P.count) fail(where + ": NORMAL count differs from POSITION"); }
            std::vector<Vec3> uv(P.count);
            if (attr.has("TEXCOORD_0")) { const Accessor T = readAccessor(D, attr.integer("TEXCOORD_0", 0)); for (size_t i = 0; i < T.count && i < P.count; ++i) uv[i] = Vec3((float)T.v[T.width * i], 1.0f - (float)T.v[T.width * i + 1], 0.0f); }
            importerVertexOrder(M.vertices, M.triangles, normals, uv);
            MeshInfo info{};
            info.vertexOffset = info.uvOffset = info.normalOffset = (int)(S.vertices.size() * 3);
            info.indexOffset = (int)(S.triangles.size() * 3);
            info.materialIndex = defaultMaterial;
            if (prim.has("material")) for (size_t r = 0; r < materialOrder.size(); ++r) if (materialOrder[r] == (size_t)prim.integer("material", 0)) info.materialIndex = (int)r;
            S.meshInfoPerMesh.push_back(info);
            S.vertices.insert(S.vertices.end(), M.vertices.begin(), M.vertices.end());
            S.triangles.insert(S.triangles.end(), M.triangles.begin(), M.triangles.end());
            S.normals.insert(S.normals.end(), normals.begin(), normals.end());
            S.uvs.insert(S.uvs.end(), uv.begin(), uv.end());
            S.meshes.push_back(std::move(M));
        }
        meshEnd[mi] = S.meshes.size();
    }
    if (S.meshes.empty()) fail(path + ": no meshes");
    // instances: depth-first, a node's primitives before its children
    const Json* nodes = doc.find("nodes");
    struct Walker
    {
        Parsed& S; const Json* nodes; const std::vector<size_t>& meshFirst; const std::vector<size_t>& meshEnd;
        void visit(size_t index, const Mat4x4& parent, std::vector<size_t>& trail)
        {
            for (size_t t : trail) if (t == index) fail("node " + std::to_string(index) + ": cycle in the node hierarchy");
            const Json& node = nodes->at(index);
            const Mat4x4 current = multiply(parent, nodeTransform(node));
            if (node.has("mesh")) {
                const size_t m = (size_t)node.integer("mesh", 0);
                if (m >= meshFirst.size()) fail("node " + std::to_string(index) + ": mesh index out of range");
                for (size_t k = meshFirst[m]; k < meshEnd[m]; ++k)
                    S.instances.push_back(Parsed::InstanceData{(unsigned)k, current, 0u, (unsigned)S.meshInfoPerMesh[k].materialIndex});
            }
            trail.push_back(index);
            if (const Json* ch = node.find("children")) for (size_t k = 0; k < ch->size(); ++k) visit((size_t)ch->at(k).number, current, trail);
            trail.pop_back();
        }
    } walker{S, nodes, meshFirst, meshEnd};
    const Json* scenes = doc.find("scenes");
    if (scenes && scenes->size() && nodes) {
        const Json& scene = scenes->at((size_t)doc.integer("scene", 0));
        std::vector<size_t> trail;
        if (const Json* roots = scene.find("nodes")) for (size_t k = 0; k < roots->size(); ++k) walker.visit((size_t)roots->at(k).number, Mat4x4(), trail);
    }
    if (S.instances.empty()) fail(path + ": the default scene instantiates no mesh");
    for (const auto& inst : S.instances) S.meshInfo.push_back(S.meshInfoPerMesh[inst.mesh]);
    return S;
}

}  // namespace gltf

#define TEX_DIM 4096   // tools/sceneBuilder.cpp:22

struct Scene
{
    // tools/sceneBuilder.h:34-48
    Buffer meshInfoData = nullptr, vertexData = nullptr, indexData = nullptr, uvData = nullptr, normalData = nullptr, materialData = nullptr;
    ImageArray textureData = nullptr;
    Sampler sampler = nullptr;
    TopAccelStruct topAccelStruct = nullptr;
    gltf::Parsed parsed;                                  // host arrays, kept for tools and tests

    // optional decoder for embedded images that are not 8-bit PNG: (bytes, size, &rgba, &w, &h) -> true when decoded
    typedef bool (*ImageDecoder)(const unsigned char*, size_t, std::vector<unsigned char>&, int&, int&);
    static ImageDecoder& decodeImage() { static ImageDecoder d = nullptr; return d; }
    static unsigned& textureDim() { static unsigned d = TEX_DIM; return d; }

    static Scene* Load(std::string path, Platform* plt, bool loadFromCache = false)
    {
        Scene* s = new Scene();
        s->parsed = gltf::parse(path);
        gltf::Parsed& P = s->parsed;
        const unsigned dim = textureDim();
        s->textureData = CreateImageArray(plt, dim, dim, P.images.empty() ? 1u : (unsigned)P.images.size());
        s->sampler = CreateSampler(plt, RD_ADDRESS_REPEAT, RD_FILTER_LINEAR);
        for (size_t k = 0; k < P.images.size(); ++k) {
            std::vector<unsigned char> rgba; int w = 0, h = 0;
            if (!gltf::decodePng(P.images[k].data(), P.images[k].size(), rgba, w, h) &&
                !(decodeImage() && decodeImage()(P.images[k].data(), P.images[k].size(), rgba, w, h)))
                gltf::fail("embedded image " + std::to_string(k) + " is not a PNG and no RD::Scene::decodeImage() is set");
            std::vector<unsigned char> layer = gltf::resizeRgba(rgba, w, h, (int)dim);
            WriteImage(plt, s->textureData, dim, dim, k, layer.data());
        }
        auto upload = [&](const void* data, size_t bytes) {
            Buffer b = CreateBuffer(plt, (unsigned)(bytes ? bytes : 16));
            if (bytes) WriteBuffer(plt, b, bytes, const_cast<void*>(data));
            return b;
        };
        s->meshInfoData = upload(P.meshInfo.data(), P.meshInfo.size() * sizeof(MeshInfo));
        s->vertexData = upload(P.vertices.data(), P.vertices.size() * sizeof(Vec3));
        s->indexData = upload(P.triangles.data(), P.triangles.size() * sizeof(Triangle));
        s->uvData = upload(P.uvs.data(), P.uvs.size() * sizeof(Vec3));
        s->normalData = upload(P.normals.data(), P.normals.size() * sizeof(Vec3));
        s->materialData = upload(P.materials.data(), P.materials.size() * sizeof(Material));
        const std::string cache = path + ".cache";
        if (loadFromCache) FileToTopAccelStruct(plt, cache.c_str(), &s->topAccelStruct);
        else {
            std::vector<BottomAccelStruct> blas;
            for (auto& m : P.meshes) { Mesh mesh; mesh.vertexData = m.vertices; mesh.indexData = m.triangles; blas.push_back(BuildAccelStruct(plt, mesh)); }
            std::vector<Instance> instances;
            for (auto& i : P.instances) instances.push_back(Instance{i.transform, i.SBTOffset, i.customInstanceID, blas[i.mesh]});
            s->topAccelStruct = BuildAccelStruct(plt, instances);
            TopAccelStructToFile(plt, s->topAccelStruct, cache.c_str());
        }
        return s;
    }
};

}  // namespace RD

#endif
