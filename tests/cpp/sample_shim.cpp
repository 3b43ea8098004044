// sample_shim.cpp -- a host program written against the reference's C++ API (namespace RD, the flow of
// samples/sample1.cpp:363-443 and its render loop :466-490) compiled against include/radiance.h.
// usage: sample_shim <program.cl> <sbt.json> <out.rgba> [width height]
// Scene: a floor quad and a box, one instance each (MeshInfo is indexed by instanceIndex), one material per
// mesh, one directional light; descriptors in the order of samples/shader.cl:176-190.
#include <cstdlib>
#include <cstdio>
#include <string>
#include <vector>
#include "radiance.h"

static void addQuad(RD::Mesh& m, RD::Vec3 a, RD::Vec3 b, RD::Vec3 c, RD::Vec3 d)
{
    const unsigned base = (unsigned)m.vertexData.size();
    m.vertexData.push_back(a); m.vertexData.push_back(b); m.vertexData.push_back(c); m.vertexData.push_back(d);
    m.indexData.push_back({base, base + 1, base + 2});
    m.indexData.push_back({base, base + 2, base + 3});
}

int main(int argc, char** argv)
{
    if (argc < 4) { fprintf(stderr, "usage: %s program.cl sbt.json out.rgba [w h]\n", argv[0]); return 2; }
    const unsigned width = argc > 5 ? (unsigned)atoi(argv[4]) : 160, height = argc > 5 ? (unsigned)atoi(argv[5]) : 120;
    RD::Platform* plt = RD::Platform::GetPlatform();

    RD::Mesh floor, box;
    addQuad(floor, {-1, 0, -1}, {1, 0, -1}, {1, 0, 1}, {-1, 0, 1});
    const float l = -0.3f, h = 0.3f, t = 0.6f;
    addQuad(box, {l, 0, l}, {l, t, l}, {h, t, l}, {h, 0, l});
    addQuad(box, {l, 0, h}, {h, 0, h}, {h, t, h}, {l, t, h});
    addQuad(box, {l, 0, l}, {l, 0, h}, {l, t, h}, {l, t, l});
    addQuad(box, {h, 0, l}, {h, t, l}, {h, t, h}, {h, 0, h});
    addQuad(box, {l, t, l}, {l, t, h}, {h, t, h}, {h, t, l});

    RD::BottomAccelStruct floorAS = RD::BuildAccelStruct(plt, floor), boxAS = RD::BuildAccelStruct(plt, box);
    std::vector<RD::Instance> instances(2);
    instances[0] = {RD::Mat4x4(), 0, 0, floorAS};
    RD::Mat4x4 xf; xf.a4 = 0.2f; xf.c4 = -0.1f;   // translate the box
    instances[1] = {xf, 0, 1, boxAS};
    RD::TopAccelStruct topAS = RD::BuildAccelStruct(plt, instances);

    // scene bundle (tools/sceneBuilder.h:8-17 layout; offsets in scalar elements)
    std::vector<float> vertexData; std::vector<unsigned> indexData;
    RD::MeshInfo info[2] = {};
    const RD::Mesh* meshes[2] = {&floor, &box};
    for (int k = 0; k < 2; ++k) {
        info[k].vertexOffset = info[k].uvOffset = info[k].normalOffset = (int)vertexData.size();
        info[k].indexOffset = (int)indexData.size();
        info[k].materialIndex = k;
        for (auto& v : meshes[k]->vertexData) { vertexData.push_back(v.x); vertexData.push_back(v.y); vertexData.push_back(v.z); }
        for (auto& f : meshes[k]->indexData) { indexData.push_back(f.idx0); indexData.push_back(f.idx1); indexData.push_back(f.idx2); }
    }
    std::vector<float> zeros(vertexData.size(), 0.0f);
    RD::Material mats[2] = {{{0.7f, 0.7f, 0.7f, 1}, 0, 0.8f, 0, 1.5f, -1, -1, -1, -1}, {{0.8f, 0.3f, 0.2f, 1}, 0, 0.5f, 0, 1.5f, -1, -1, -1, -1}};
    RD::SceneProperties scene = {};
    scene.lightCount[0] = 1;
    const float dir[4] = {-0.4f, -1.0f, -0.3f, 0}, col[4] = {10, 10, 10, 1};
    memcpy(scene.lights[0].direction, dir, 16); memcpy(scene.lights[0].color, col, 16);
    RD::PhysicalCamera cam = {(float)width, (float)height, 0.030f, 0.036f, 2.0f, 0.0f, 0.0f, 0.8f, 2.6f, -0.2f, 0.0f, 0.0f};
    RD::RayTraceProperties RTProp = {0, 2, 1, 0};

    auto upload = [&](void* data, size_t bytes) { RD::Buffer b = RD::CreateBuffer(plt, (unsigned)bytes); RD::WriteBuffer(plt, b, bytes, data); return b; };
    RD::Buffer rdRTProp = upload(&RTProp, sizeof(RTProp));
    RD::Buffer rdImageScratch = RD::CreateBuffer(plt, width * height * RD_CHANNEL * sizeof(float));
    RD::Image rdImage = RD::CreateImage(plt, width, height);
    RD::Buffer rdCam = upload(&cam, sizeof(cam)), rdScene = upload(&scene, sizeof(scene)), rdMeshInfo = upload(info, sizeof(info));
    RD::Buffer rdVertex = upload(vertexData.data(), vertexData.size() * 4), rdIndex = upload(indexData.data(), indexData.size() * 4);
    RD::Buffer rdUV = upload(zeros.data(), zeros.size() * 4), rdNormal = upload(zeros.data(), zeros.size() * 4);
    RD::Buffer rdMaterial = upload(mats, sizeof(mats));
    RD::ImageArray rdTextures = RD::CreateImageArray(plt, 4, 4, 1);
    RD::Sampler rdSampler = RD::CreateSampler(plt, RD_ADDRESS_REPEAT, RD_FILTER_LINEAR);

    RD::DescriptorSet descSet = RD::CreateDescriptorSet({rdRTProp, rdImageScratch, rdImage, rdCam, rdScene, rdMeshInfo, rdVertex,
                                                         rdIndex, rdUV, rdNormal, rdMaterial, rdTextures, rdSampler, topAS});
    RD::PipelineLayout layout = RD::CreatePipelineLayout({RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::IMAGE_TYPE, RD::BUFFER_TYPE, RD::BUFFER_TYPE,
        RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::BUFFER_TYPE, RD::TEX_ARRAY_TYPE,
        RD::IMAGE_SAMPLER_TYPE, RD::ACCEL_STRUCT_TYPE});

    char *code = nullptr, *sbt = nullptr; size_t codeSize = 0, sbtSize = 0;
    if (RD::read_kernel_file_str(argv[1], &code, &codeSize) || RD::read_kernel_file_str(argv[2], &sbt, &sbtSize)) return 3;
    RD::LoadShaderBindingTable(plt, sbt);   // extension: replaces the offline tools/genSBT.py step
    // RD_SAMPLE_STREAM=1: the same loop with the B200 extension that copies the frame out band by band while it renders
    const bool streamed = getenv("RD_SAMPLE_STREAM") != nullptr;
    if (streamed && rd_set_option(plt->impl, "stream_image", 1)) return 5;
    RD::ShaderModule module = RD::CreateShaderModule(plt, code, (unsigned)codeSize, (char*)"raygen");
    RD::Pipeline pipeline = RD::CreatePipeline({2, layout, {module}, {}});
    RD::BindPipeline(plt, pipeline);
    RD::BindDescriptorSet(plt, descSet);

    // progressive loop of samples/sample1.cpp:466-490: two calls of two samples each
    std::vector<unsigned char> image(width * height * RD_CHANNEL);
    if (streamed) RD::StreamBufferToHost(plt, rdImage, image.data(), RD_CHANNEL, 8);
    for (int call = 0; call < 2; ++call) {
        RD::WriteBuffer(plt, rdRTProp, sizeof(RTProp), &RTProp);
        RD::TraceRays(plt, 0, 0, 0, width, height);
        RTProp.totalSamples += RTProp.batchSize;
        RD::ReadBuffer(plt, rdImage, image.size(), image.data());
    }
    FILE* f = fopen(argv[3], "wb");
    if (!f) return 4;
    fwrite(image.data(), 1, image.size(), f);
    fclose(f);
    unsigned long long sum = 0;
    for (unsigned char c : image) sum += c;
    printf("sample_shim: %ux%u, checksum %llu\n", width, height, sum);
    return 0;
}
