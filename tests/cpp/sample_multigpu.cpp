// sample_multigpu.cpp -- a C++ host that renders ONE frame on all GPUs of the box through the C ABI alone
// (include/radiance_b200.h): one host thread per GPU, the scene replicated, interleaved bands of 8x4-tile rows per rank,
// the frame gathered on rank 0 by rd_frame_gather (NCCL inside the library). No Python, no torch: this is what a host in
// the reference's idiom (samples/sample1.cpp:363-490, one device there) adds to go multi-GPU.
// usage: sample_multigpu <program.cl> <sbt.json> <gpus> [width height]   -- exit code 0 iff the gathered frame equals the
// frame rank 0 renders alone, for the static partition and for the dynamic work-unit partition.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "radiance_b200.h"

#define CHECK(call) do { int _rc = (call); if (_rc != RD_OK) { fprintf(stderr, "rank %d: %s failed (%d): %s\n", rank, #call, _rc, rd_last_error()); ok[rank] = 0; return; } } while (0)

static std::string readFile(const char* path)
{
    std::string s; FILE* f = fopen(path, "rb");
    if (!f) return s;
    char buf[4096]; size_t n;
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) s.append(buf, n);
    fclose(f);
    return s;
}

struct MeshInfo { int vertexOffset, indexOffset, uvOffset, normalOffset, materialIndex, pad[3]; };
struct Material { float albedo[4], metallic, roughness, transmission, ior; int albedoTex, metallicTex, roughnessTex, normalTex; };
struct Scene { unsigned lightCount[4]; struct { float direction[4], color[4]; } lights[5]; };
struct Camera { float widthPixel, heightPixel, focalLength, sensorWidth, focalDistance, fStop, x, y, z, wx, wy, wz; };

static std::mutex g_compile;
static unsigned char g_uid[128];
static rd_platform* g_owner = 0;
static std::vector<int> ok;
static std::vector<unsigned char> g_whole, g_static, g_dynamic;

static void addQuad(std::vector<float>& v, std::vector<unsigned>& idx, const float (&p)[4][3])
{
    const unsigned base = (unsigned)v.size() / 3;
    for (int k = 0; k < 4; ++k) { v.push_back(p[k][0]); v.push_back(p[k][1]); v.push_back(p[k][2]); }
    const unsigned t[6] = {base, base + 1, base + 2, base, base + 2, base + 3};
    idx.insert(idx.end(), t, t + 6);
}

static void rankMain(int rank, int world, const std::string& code, const std::string& sbt, unsigned w, unsigned h,
                     std::vector<std::thread>* barrierDummy)
{
    (void)barrierDummy;
    rd_platform* p = 0;
    CHECK(rd_platform_create(rank, &p));
    // floor + box, one instance each
    std::vector<float> fv, bv; std::vector<unsigned> fi, bi;
    const float F[4][3] = {{-1, 0, -1}, {1, 0, -1}, {1, 0, 1}, {-1, 0, 1}};
    addQuad(fv, fi, F);
    const float l = -0.3f, hh = 0.3f, t = 0.6f;
    const float B[5][4][3] = {{{l, 0, l}, {l, t, l}, {hh, t, l}, {hh, 0, l}}, {{l, 0, hh}, {hh, 0, hh}, {hh, t, hh}, {l, t, hh}},
                              {{l, 0, l}, {l, 0, hh}, {l, t, hh}, {l, t, l}}, {{hh, 0, l}, {hh, t, l}, {hh, t, hh}, {hh, 0, hh}},
                              {{l, t, l}, {l, t, hh}, {hh, t, hh}, {hh, t, l}}};
    for (auto& q : B) addQuad(bv, bi, q);
    rd_blas *fa = 0, *ba = 0;
    CHECK(rd_blas_build(p, fv.data(), (unsigned)fv.size() / 3, fi.data(), (unsigned)fi.size() / 3, &fa));
    CHECK(rd_blas_build(p, bv.data(), (unsigned)bv.size() / 3, bi.data(), (unsigned)bi.size() / 3, &ba));
    rd_instance_desc inst[2];
    memset(inst, 0, sizeof(inst));
    for (int k = 0; k < 2; ++k) { inst[k].transform[0] = inst[k].transform[5] = inst[k].transform[10] = inst[k].transform[15] = 1.0f; }
    inst[1].transform[3] = 0.2f; inst[1].transform[11] = -0.1f;
    inst[0].bottomAccelStruct = fa; inst[1].bottomAccelStruct = ba; inst[1].customInstanceID = 1;
    rd_handle tlas = 0;
    CHECK(rd_tlas_build(p, inst, 2, &tlas));
    std::vector<float> verts(fv); verts.insert(verts.end(), bv.begin(), bv.end());
    std::vector<unsigned> index(fi); index.insert(index.end(), bi.begin(), bi.end());
    MeshInfo info[2]; memset(info, 0, sizeof(info));
    info[1].vertexOffset = info[1].uvOffset = info[1].normalOffset = (int)fv.size(); info[1].indexOffset = (int)fi.size(); info[1].materialIndex = 1;
    Material mats[2] = {{{0.7f, 0.7f, 0.7f, 1}, 0, 0.8f, 0, 1.5f, -1, -1, -1, -1}, {{0.8f, 0.3f, 0.2f, 1}, 0, 0.5f, 0, 1.5f, -1, -1, -1, -1}};
    Scene scene; memset(&scene, 0, sizeof(scene));
    scene.lightCount[0] = 1;
    const float dir[4] = {-0.4f, -1.0f, -0.3f, 0}, col[4] = {10, 10, 10, 1};
    memcpy(scene.lights[0].direction, dir, 16); memcpy(scene.lights[0].color, col, 16);
    Camera cam = {(float)w, (float)h, 0.030f, 0.036f, 2.0f, 0.0f, 0.0f, 0.8f, 2.6f, -0.2f, 0.0f, 0.0f};
    unsigned rt[4] = {0, 4, 4, 0};
    std::vector<float> zeros(verts.size(), 0.0f);
    auto upload = [&](const void* d, size_t n, rd_handle* out) { int rc = rd_buffer_create(p, n, out); return rc ? rc : rd_buffer_write(p, *out, n, d, 0); };
    rd_handle hRt, hScratch, hImage, hCam, hScene, hInfo, hV, hI, hUV, hN, hM, hTex, hSmp;
    CHECK(upload(rt, 16, &hRt)); CHECK(rd_buffer_create(p, (size_t)w * h * 16, &hScratch)); CHECK(rd_image_create(p, w, h, &hImage));
    CHECK(upload(&cam, sizeof(cam), &hCam)); CHECK(upload(&scene, sizeof(scene), &hScene)); CHECK(upload(info, sizeof(info), &hInfo));
    CHECK(upload(verts.data(), verts.size() * 4, &hV)); CHECK(upload(index.data(), index.size() * 4, &hI));
    CHECK(upload(zeros.data(), zeros.size() * 4, &hUV)); CHECK(upload(zeros.data(), zeros.size() * 4, &hN)); CHECK(upload(mats, sizeof(mats), &hM));
    CHECK(rd_image_array_create(p, 4, 4, 1, &hTex)); CHECK(rd_sampler_create(p, 0x1133, 0x1141, &hSmp));
    rd_handle desc[14] = {hRt, hScratch, hImage, hCam, hScene, hInfo, hV, hI, hUV, hN, hM, hTex, hSmp, tlas};
    rd_module* mod = 0;
    {
        std::lock_guard<std::mutex> lock(g_compile);
        CHECK(rd_sbt_load_json(p, sbt.c_str()));
        CHECK(rd_shader_module_create(p, code.c_str(), (unsigned)code.size(), "raygen", &mod));
    }
    CHECK(rd_pipeline_bind(p, &mod, 1));
    CHECK(rd_descriptors_bind(p, desc, 14));
    std::vector<unsigned char> zero((size_t)w * h * 4, 0);
    if (rank == 0) {   // the frame one GPU renders alone
        CHECK(rd_trace_rays(p, 0, 0, 0, w, h));
        g_whole.resize(zero.size());
        CHECK(rd_buffer_read(p, hImage, g_whole.size(), g_whole.data(), 0));
    }
    CHECK(rd_comm_init(p, rank, world, g_uid));
    // the job-wide unit counter lives on rank 0's GPU. Ranks in separate PROCESSES exchange the 64-byte IPC handle
    // (rd_steal_counter_export / _import, tools/exp_multigpu.py); host threads of one process adopt it through peer access
    if (rank == 0) { unsigned char handle[64]; CHECK(rd_steal_counter_export(p, handle)); g_owner = p; }
    static std::mutex m; static int arrived = 0;
    { std::lock_guard<std::mutex> lk(m); ++arrived; }
    while (true) { std::lock_guard<std::mutex> lk(m); if (arrived >= world) break; }
    if (rank != 0) CHECK(rd_steal_counter_adopt(p, g_owner));
    // static partition: interleaved bands + one packed send per peer
    CHECK(rd_buffer_write(p, hImage, zero.size(), zero.data(), 0));
    CHECK(rd_trace_rays_interleaved(p, w, h, (unsigned)rank, (unsigned)world, 1));
    CHECK(rd_frame_gather(p, hImage, w, h, 4, 1));
    if (rank == 0) { g_static.resize(zero.size()); CHECK(rd_buffer_read(p, hImage, g_static.size(), g_static.data(), 0)); }
    else CHECK(rd_platform_sync(p));
    // dynamic partition: work units from the job-wide counter
    CHECK(rd_buffer_write(p, hImage, zero.size(), zero.data(), 0));
    unsigned long long owned = 0;
    CHECK(rd_trace_rays_units(p, w, h, 4u * (unsigned)world, 7, &owned));
    CHECK(rd_frame_gather_units(p, hImage, w, h, 4, 4u * (unsigned)world, owned));
    if (rank == 0) { g_dynamic.resize(zero.size()); CHECK(rd_buffer_read(p, hImage, g_dynamic.size(), g_dynamic.data(), 0)); }
    else CHECK(rd_platform_sync(p));
    printf("rank %d of %d: rendered %d of %d units\n", rank, world, __builtin_popcountll(owned), 4 * world);
    CHECK(rd_comm_destroy(p));
    rd_module_free(p, mod);
    rd_platform_destroy(p);
}

int main(int argc, char** argv)
{
    if (argc < 4) { fprintf(stderr, "usage: %s program.cl sbt.json gpus [w h]\n", argv[0]); return 2; }
    const std::string code = readFile(argv[1]), sbt = readFile(argv[2]);
    const int world = atoi(argv[3]);
    const unsigned w = argc > 5 ? (unsigned)atoi(argv[4]) : 320, h = argc > 5 ? (unsigned)atoi(argv[5]) : 240;
    if (code.empty() || sbt.empty() || world < 1) return 3;
    if (rd_comm_unique_id(g_uid) != RD_OK) { fprintf(stderr, "%s\n", rd_last_error()); return 4; }
    ok.assign((size_t)world, 1);
    std::vector<std::thread> th;
    for (int r = 0; r < world; ++r) th.emplace_back(rankMain, r, world, std::cref(code), std::cref(sbt), w, h, &th);
    for (auto& t : th) t.join();
    for (int r = 0; r < world; ++r) if (!ok[(size_t)r]) return 5;
    const bool a = g_static == g_whole, b = g_dynamic == g_whole;
    unsigned long long sum = 0;
    for (unsigned char c : g_whole) sum += c;
    printf("sample_multigpu: %d GPU(s), %ux%u, checksum %llu, static partition %s, dynamic partition %s\n", world, w, h, sum,
           a ? "identical" : "DIFFERENT", b ? "identical" : "DIFFERENT");
    return a && b && sum > 0 ? 0 : 1;
}
