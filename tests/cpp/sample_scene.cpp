// sample_scene.cpp -- the asset path of samples/sample1.cpp (RD::Scene::Load :375, INCLUDE_SCENE_DESC / _LAYOUT :387-407,
// the `render` callback :447-499 driven by renderLoop :440) written against include/radiance.h, include/sceneBuilder.h and
// include/inspector.h.
// usage: sample_scene <scene.glb> <program.cl> <sbt.json> <out prefix> [texture dim] [width height]
// Writes <prefix>.arrays (the descriptor arrays the loader produced), <prefix>_<frame>.png via renderLoop and
// <prefix>.rgba (the last image). Loads the scene twice: built, then from the <scene>.cache blob the first load wrote.
#include <cstdio>
#include <string>
#include <vector>
#include "sceneBuilder.h"
#include "inspector.h"

struct CbData   // samples/sample1.cpp:344-361
{
    RD::Platform* plt;
    unsigned extent[2];
    RD::Image rdImage; RD::Buffer rdRTProp;
    std::vector<unsigned char> image;
    int frames = 0;
};

static void render(void* data, unsigned char** image, int* out_width, int* out_height)
{
    CbData* d = (CbData*)data;
    RD::TraceRays(d->plt, 0, 0, 0, d->extent[0], d->extent[1]);
    RD::ReadBuffer(d->plt, d->rdImage, d->image.size(), d->image.data());
    RD::RayTraceProperties RTProp;
    RD::ReadBuffer(d->plt, d->rdRTProp, sizeof(RTProp), &RTProp);
    RTProp.totalSamples += RTProp.batchSize;
    RD::WriteBuffer(d->plt, d->rdRTProp, sizeof(RTProp), &RTProp);
    d->frames++;
    *image = d->image.data(); *out_width = (int)d->extent[0]; *out_height = (int)d->extent[1];
}

template <class T> static void dump(FILE* f, const std::vector<T>& v)
{
    const unsigned long long n = v.size() * sizeof(T);
    fwrite(&n, 8, 1, f);
    if (n) fwrite(v.data(), 1, n, f);
}

int main(int argc, char** argv)
{
    if (argc < 5) { fprintf(stderr, "usage: %s scene.glb program.cl sbt.json prefix [texdim] [w h]\n", argv[0]); return 2; }
    const std::string prefix = argv[4];
    if (argc > 5) RD::Scene::textureDim() = (unsigned)atoi(argv[5]);
    const unsigned width = argc > 7 ? (unsigned)atoi(argv[6]) : 320, height = argc > 7 ? (unsigned)atoi(argv[7]) : 240;
    RD::Platform* plt = RD::Platform::GetPlatform();
    RD::Scene* scene = RD::Scene::Load(argv[1], plt);
    RD::Scene* cached = RD::Scene::Load(argv[1], plt, true);

    {   // the arrays behind the nine descriptors, for the test to compare with the Python loader
        FILE* f = fopen((prefix + ".arrays").c_str(), "wb");
        if (!f) return 4;
        const RD::gltf::Parsed& P = scene->parsed;
        dump(f, P.meshInfo); dump(f, P.vertices); dump(f, P.triangles); dump(f, P.uvs); dump(f, P.normals); dump(f, P.materials);
        std::vector<float> inst;
        for (auto& i : P.instances) { inst.push_back((float)i.mesh); inst.push_back((float)i.customInstanceID); for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) inst.push_back(i.transform[r][c]); }
        dump(f, inst);
        fclose(f);
    }

    RD::RayTraceProperties RTProp = {0, 1, 1, 0};
    RD::PhysicalCamera cam = {(float)width, (float)height, 0.030f, 0.036f, 2.0f, 0.0f, 0.02f, 0.9f, 2.2f, -0.3f, 0.0f, 0.0f};
    RD::SceneProperties light = {};
    light.lightCount[0] = 1;
    const float dir[4] = {-0.4f, -1.0f, -0.3f, 0}, col[4] = {10, 10, 10, 1};
    memcpy(light.lights[0].direction, dir, 16); memcpy(light.lights[0].color, col, 16);
    auto upload = [&](void* data, size_t bytes) { RD::Buffer b = RD::CreateBuffer(plt, (unsigned)bytes); RD::WriteBuffer(plt, b, bytes, data); return b; };
    RD::Buffer rdRTProp = upload(&RTProp, sizeof(RTProp));
    RD::Buffer rdImageScratch = RD::CreateBuffer(plt, width * height * RD_CHANNEL * sizeof(float));
    RD::Image rdImage = RD::CreateImage(plt, width, height);
    RD::Buffer rdCam = upload(&cam, sizeof(cam)), rdLight = upload(&light, sizeof(light));

    RD::PipelineLayout layout = RD::CreatePipelineLayout({RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::IMAGE_TYPE, RD::BUFFER_TYPE, RD::BUFFER_TYPE, INCLUDE_SCENE_LAYOUT});
    char *code = nullptr, *sbt = nullptr; size_t codeSize = 0, sbtSize = 0;
    if (RD::read_kernel_file_str(argv[2], &code, &codeSize) || RD::read_kernel_file_str(argv[3], &sbt, &sbtSize)) return 3;
    RD::LoadShaderBindingTable(plt, sbt);
    RD::ShaderModule module = RD::CreateShaderModule(plt, code, (unsigned)codeSize, (char*)"raygen");
    RD::Pipeline pipeline = RD::CreatePipeline({2, layout, {module}, {}});
    RD::BindPipeline(plt, pipeline);

    CbData data{plt, {width, height}, rdImage, rdRTProp, std::vector<unsigned char>(width * height * RD_CHANNEL)};
    std::vector<unsigned char> first;
    for (RD::Scene* s : {scene, cached}) {
        RD::DescriptorSet descSet = RD::CreateDescriptorSet({rdRTProp, rdImageScratch, rdImage, rdCam, rdLight, INCLUDE_SCENE_DESC(s)});
        RD::BindDescriptorSet(plt, descSet);
        RTProp.totalSamples = 0;
        RD::WriteBuffer(plt, rdRTProp, sizeof(RTProp), &RTProp);
        RD::inspector::config().frames = 2;
        RD::inspector::config().output = s == scene ? prefix + "_%d.png" : std::string();
        if (renderLoop(render, &data)) return 5;
        if (s == scene) first = data.image;
        else if (first != data.image) { printf("sample_scene: the cached structure renders a different image\n"); return 6; }
    }
    FILE* f = fopen((prefix + ".rgba").c_str(), "wb");
    if (!f) return 4;
    fwrite(data.image.data(), 1, data.image.size(), f);
    fclose(f);
    printf("sample_scene: %zu meshes, %zu instances, %zu images, %d frames\n", scene->parsed.meshes.size(), scene->parsed.instances.size(),
           scene->parsed.images.size(), data.frames);
    return 0;
}
